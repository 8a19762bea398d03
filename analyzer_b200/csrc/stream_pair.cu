// Streaming (dataflow) rating kernel for the reference-faithful layout: every
// player carries the shared pair (player.trueskill_mu/_sigma) AND the pair of the
// one queue this ladder is rated in (player.trueskill_<mode>_mu/_sigma), and every
// match does what rater.py:108-169 does -- quality on the queue pairs, rate(shared)
// with trueskill_delta, rate(queue), NULL fallbacks (shared NULL -> seed, queue
// NULL -> the shared pair just read) -- in stream order with the version protocol
// of stream.cu instead of rounds and barriers.
//
// Row format (workspace, 64 bytes = two sectors per player): eight 64-bit words,
//   shared mu lo|ver, mu hi|ver, sigma lo|ver, sigma hi|ver, queue mu lo|ver, ...
// all eight carry the number of the last match that wrote the row; both pairs are
// always read and written together, so one version covers them.
//
// MULTI: the batch rates matches of several game modes against a table with several
// queue columns (the faithful 8-pair row; worker.py:176-192 rates all modes in one
// chronological stream).  A player's shared pair is rewritten by every match of the
// player, a queue pair only by the matches of that queue, so a queue pair cannot carry
// the number of the player's previous match.  Instead every queue pair carries an
// ORDINAL -- how many matches of that queue have written it -- and the ordinals of all
// of a player's queues travel with the shared pair:
//   words 0..3   shared pair, version = last match of the player            (as above)
//   words 4..5   ordinals of queues 0..2 | 3..5, 10 bits each, same version
//   words 8+4j.. pair of queue j (compact index over the rated columns), version =
//                its ordinal
// A slot reads the three pieces in one poll: shared and ordinals must carry pred, the
// queue pair must carry the ordinal the second word names.  The match then writes the
// queue pair under ordinal + 1, and shared pair and ordinals under its own number.
// No second link array, no fence: every 8-byte word still validates itself.  The queue
// pair's 10-bit tag distinguishes ordinal k from k - 1 (and anything less than 1024
// writes older); a load at gpu scope cannot be staler than that, L2 holds the one copy.
#include "stream_common.cuh"

namespace {

using namespace tsstream;
using tsvw::trunc_vw;

constexpr int kBlock = 256;
// 1: issue the loads of all missing slots before checking any (measured slower: config 2, K = 2,
// rate 5.16 vs 3.75 ms -- more loads in flight per poll cost more than the round trips they save;
// polling only the first missing slot per iteration is slower too, 3.93 ms)
#ifndef TS_PAIR_HOIST
#define TS_PAIR_HOIST 0
#endif

struct PairArgs {
    ts_params prm;
    unsigned long long *vt;       // row_words words per player
    const double *seed;
    ModeCols modes;               // >= 0: the mode is rated
    const uint32_t *ids;
    const uint32_t *meta;
    const uint32_t *pred;
    uint32_t row_words;           // words per player row: 8, MULTI: 8 + 4 * (rated queue columns)
    uint32_t queue_nibbles;       // MULTI: 4 bits per game mode, the compact index of its queue column
    uint32_t n_matches;
    uint32_t n_players;           // ids at or beyond it are refused, never dereferenced
    uint32_t seq_base;
    uint32_t sleep_ns;
    uint32_t max_spins;
    uint32_t prefetch;            // 1: the rows of the next tile's players are prefetched into L2
    ts_outputs out;
    unsigned long long *ctrl;
};

constexpr uint32_t kOrdMask = 1023u;

__device__ __forceinline__ void ld_ord(const unsigned long long *p, unsigned long long (&w)[2])
{
    asm volatile("ld.relaxed.gpu.global.v2.u64 {%0, %1}, [%2];" : "=l"(w[0]), "=l"(w[1]) : "l"(p) : "memory");
}
__device__ __forceinline__ void st_ord(unsigned long long *p, uint32_t lo, uint32_t hi, uint32_t ver)
{
    const unsigned long long v = (unsigned long long)ver << 32;
    asm volatile("st.relaxed.gpu.global.v2.u64 [%0], {%1, %2};" ::"l"(p), "l"(v | lo), "l"(v | hi) : "memory");
}

template <int P, int L, bool MULTI>
__global__ void __launch_bounds__(kBlock, 2) k_rate_stream_pair(const PairArgs a)
{
    constexpr int W = 32 / L;
    constexpr int SPL = (P + L - 1) / L;
    constexpr int HALF = P / 2;
    constexpr int kWarps = kBlock / 32;
    __shared__ __align__(16) WarpStage<P, W> stage_all[kWarps];

    const int wl = threadIdx.x & 31;
    const int lane = wl % L;
    const int grp = wl / L;
    WarpStage<P, W> &stage = stage_all[threadIdx.x >> 5];
    const double tau2 = a.prm.tau * a.prm.tau;
    const double beta2 = a.prm.beta * a.prm.beta;
    const double tb2 = tau2 + beta2;
    const double margin_full = a.prm.draw_margin[P];
    uint32_t rated_mask = 0;
#pragma unroll
    for (int i = 0; i < 6; ++i) rated_mask |= a.modes.col[i] >= 0 ? (1u << i) : 0u;
    const uint32_t n_tiles = (a.n_matches + W - 1) / W;
    const uint32_t stride = gridDim.x * kWarps;
    const uint32_t first_tile = (threadIdx.x >> 5) * gridDim.x + blockIdx.x;

    const unsigned long long pol = stream_policy();
    auto stage_tile = [&](int b, uint32_t t) {
        if (t < n_tiles) stage_warp_tile<P, W>(stage, b, a.ids, a.pred, a.meta, t * W, min((uint32_t)W, a.n_matches - t * W), wl, pol);
    };
    int cur_buf = 0;
    stage_tile(0, first_tile);
    constexpr int D = WarpStage<P, W>::kDepth;
    if (D == 3) stage_tile(1, first_tile + stride);
    for (uint32_t tile = first_tile; tile < n_tiles; tile += stride) {
        cp_async_wait_all();                    // this tile's records and the next one's (issued a whole tile ago)
        __syncwarp();
        const int cur = cur_buf, nxt = cur == D - 1 ? 0 : cur + 1;
        cur_buf = nxt;
        stage_tile((cur + D - 1) % D, tile + (D - 1) * stride);
        if (D == 3 && a.prefetch && tile + stride < n_tiles)    // next tile's rows on their way into L2 (the shared pair's sector;
            prefetch_tile_rows<P, W>(stage, nxt, a.vt, MULTI ? a.row_words : 8u,                   // the queue pair follows it)
                                     min((uint32_t)W, a.n_matches - (tile + stride) * W), a.n_players, wl);
        const uint32_t base = tile * W;
        const uint32_t cnt = min((uint32_t)W, a.n_matches - base);
        const uint32_t m = base + grp;
        const uint32_t meta = (uint32_t)grp < cnt ? stage.meta[cur][grp] : (15u << TS_META_MODE_SHIFT);
        const uint32_t mode = (meta >> TS_META_MODE_SHIFT) & 15u;
        const bool skip_mode = ((rated_mask >> mode) & 1u) == 0u;
        const bool gated = skip_mode || (meta & (TS_META_AFK | TS_META_INVALID)) != 0u;
        const bool live = (uint32_t)grp < cnt && !gated;
        if ((uint32_t)grp < cnt && gated && lane == 0) {     // rater.py:83-85,102-106: outputs only
            if (a.out.status) a.out.status[m] = skip_mode ? TS_ST_SKIP_MODE : ((meta & TS_META_INVALID) ? TS_ST_INVALID : TS_ST_AFK);
            if (!skip_mode && a.out.quality) a.out.quality[m] = 0.0;
        }
        const int rank0 = (meta & TS_META_WIN_A) ? 0 : 1, rank1 = (meta & TS_META_WIN_B) ? 0 : 1;
        const int first = rank1 < rank0 ? 1 : 0;
        const bool is_draw = rank0 == rank1;

        // this match's queue pair in a player row: compact queue index qj, word offset q_word
        const uint32_t row_words = MULTI ? a.row_words : 8u;
        const uint32_t qj = MULTI ? (a.queue_nibbles >> ((mode & 7u) * 4u)) & 7u : 0u;
        const uint32_t q_word = MULTI ? 8u + 4u * qj : 4u;
        const uint32_t ord_shift = 10u * (qj % 3u);
        uint32_t pid[SPL], want[SPL];
        uint32_t ord_lo[SPL], ord_hi[SPL];          // MULTI: the player's queue ordinals as read
        bool used[SPL], have[SPL];
        double s_mu[SPL], s_sg[SPL], q_mu[SPL], q_sg[SPL];   // the row as read (NaN = NULL)
        bool bad_id = false;
#pragma unroll
        for (int k = 0; k < SPL; ++k) {
            const int s = lane + k * L;
            pid[k] = (live && s < P) ? stage.ids[cur][grp * P + s] : TS_EMPTY_SLOT;
            used[k] = pid[k] != TS_EMPTY_SLOT;
            if (used[k] && pid[k] >= a.n_players) { used[k] = false; bad_id = true; }   // corrupt id: no table access
            want[k] = used[k] ? stage.link[cur][grp * P + s] : TS_NONE;
            ord_lo[k] = ord_hi[k] = 0u;
            have[k] = !used[k];
            s_mu[k] = 0.0; s_sg[k] = 1.0; q_mu[k] = 0.0; q_sg[k] = 1.0;
        }
        bool done = !live;
        // wait loop as in stream.cu: one ballot per poll, and the arithmetic of every pass is
        // run by the whole converged warp (lanes of incomplete groups compute on what they
        // have and throw the result away); only the stores are predicated
        uint32_t spins = 0;
        while (!__all_sync(0xFFFFFFFFu, done)) {
            bool ready = false;
            if (!done) {
                ready = true;
#if TS_PAIR_HOIST
                unsigned long long ws[SPL][4], wq[SPL][4], wo[SPL][2];
#pragma unroll
                for (int k = 0; k < SPL; ++k) {
                    if (have[k]) continue;
                    const unsigned long long *row = a.vt + (size_t)pid[k] * row_words;
                    ld_row(row, ws[k]);
                    ld_row(row + q_word, wq[k]);
                    if (MULTI) ld_ord(row + 4, wo[k]);
                }
#pragma unroll
                for (int k = 0; k < SPL; ++k) {
                    if (have[k]) continue;
                    bool ok;
                    if (MULTI) {
                        const uint32_t ords = (uint32_t)wo[k][qj >= 3u ? 1 : 0];
                        ok = row_has_version(ws[k], want[k]) && (uint32_t)(wo[k][0] >> 32) == want[k] &&
                             (uint32_t)(wo[k][1] >> 32) == want[k] && row_has_version(wq[k], (ords >> ord_shift) & kOrdMask);
                        ord_lo[k] = (uint32_t)wo[k][0]; ord_hi[k] = (uint32_t)wo[k][1];
                    } else {
                        ok = row_has_version(ws[k], want[k]) && row_has_version(wq[k], want[k]);
                    }
                    if (ok) {
                        have[k] = true;
                        s_mu[k] = row_mu(ws[k]); s_sg[k] = row_sigma(ws[k]);
                        q_mu[k] = row_mu(wq[k]); q_sg[k] = row_sigma(wq[k]);
                    }
                    ready &= have[k];
                }
#else
#pragma unroll
                for (int k = 0; k < SPL; ++k) {
                    if (have[k]) continue;
                    unsigned long long ws[4], wq[4];
                    const unsigned long long *row = a.vt + (size_t)pid[k] * row_words;
                    ld_row(row, ws);
                    ld_row(row + q_word, wq);
                    bool ok;
                    if (MULTI) {
                        unsigned long long wo[2];
                        ld_ord(row + 4, wo);
                        const uint32_t ords = (uint32_t)wo[qj >= 3u ? 1 : 0];
                        ok = row_has_version(ws, want[k]) && (uint32_t)(wo[0] >> 32) == want[k] &&
                             (uint32_t)(wo[1] >> 32) == want[k] && row_has_version(wq, (ords >> ord_shift) & kOrdMask);
                        ord_lo[k] = (uint32_t)wo[0]; ord_hi[k] = (uint32_t)wo[1];
                    } else {
                        ok = row_has_version(ws, want[k]) && row_has_version(wq, want[k]);
                    }
                    if (ok) {
                        have[k] = true;
                        s_mu[k] = row_mu(ws); s_sg[k] = row_sigma(ws);
                        q_mu[k] = row_mu(wq); q_sg[k] = row_sigma(wq);
                    }
                    ready &= have[k];
                }
#endif
            }
            const uint32_t complete = complete_groups<L>(ready);
            if (complete == 0u) {
                if ((++spins & 0x3FFu) == 0u) {      // watchdog, see stream.cu
                    if (spins >= a.max_spins && wl == 0) atomicExch(&a.ctrl[CTRL_STALLED], 1ull);
                    if (ld_relaxed_u64(&a.ctrl[CTRL_STALLED]) != 0ull) return;
                }
                __nanosleep(a.sleep_ns);
                continue;
            }
            const bool go = ((complete >> (grp * L)) & 1u) != 0u;
            // lanes that only tag along take the branch of the first complete group
            // (the shuffle is executed by every lane: it must not sit inside the conditional)
            const int lead_draw = __shfl_sync(0xFFFFFFFFu, (int)is_draw, __ffs(complete) - 1);
            const bool draw_eff = go ? is_draw : lead_draw != 0;

            double sgn[SPL], es_mu[SPL], es_sg[SPL], eq_mu[SPL], eq_sg[SPL];   // effective pairs
            bool had[SPL];
            double d_s = 0.0, c2_s = 0.0, d_q = 0.0, c2_q = 0.0, ss_q = 0.0;
            uint32_t cnts = bad_id ? 65536u : 0u;
            bool any_null = false;
#pragma unroll
            for (int k = 0; k < SPL; ++k) {
                es_mu[k] = s_mu[k]; es_sg[k] = s_sg[k]; eq_mu[k] = q_mu[k]; eq_sg[k] = q_sg[k];
                had[k] = !isnan(s_mu[k]);
                any_null |= used[k] && (!had[k] || isnan(q_mu[k]));
            }
            if (__any_sync(0xFFFFFFFFu, any_null)) {
#pragma unroll
                for (int k = 0; k < SPL; ++k) {
                    if (!used[k]) continue;
                    if (!had[k]) {                      // rater.py:118-119: first touch -> seed
                        double2 sd = make_double2(s_mu[k], s_sg[k]);
                        if (a.seed) sd = __ldg(reinterpret_cast<const double2 *>(a.seed) + pid[k]);
                        if (isnan(sd.x)) cnts |= 65536u;    // neither rating nor seed
                        es_mu[k] = sd.x; es_sg[k] = sd.y;
                    }
                    if (isnan(q_mu[k])) { eq_mu[k] = es_mu[k]; eq_sg[k] = es_sg[k]; }   // rater.py:127-130
                }
            }
#pragma unroll
            for (int k = 0; k < SPL; ++k) {
                const int s = lane + k * L;
                sgn[k] = ((s >= HALF ? 1 : 0) == first) ? 1.0 : -1.0;
                const bool u = used[k];
                const bool bad = u && ((eq_sg[k] == 0.0) || (es_sg[k] == 0.0));
                d_s = fma(sgn[k], u ? es_mu[k] : 0.0, d_s);
                c2_s += u ? fma(es_sg[k], es_sg[k], tb2) : 0.0;
                d_q = fma(sgn[k], u ? eq_mu[k] : 0.0, d_q);
                c2_q += u ? fma(eq_sg[k], eq_sg[k], tb2) : 0.0;
                ss_q = u ? fma(eq_sg[k], eq_sg[k], ss_q) : ss_q;
                cnts += (u ? (sgn[k] > 0.0 ? 1u : 256u) : 0u) + (bad ? 65536u : 0u);
            }
            d_s = group_sum<L>(d_s);
            c2_s = group_sum<L>(c2_s);
            d_q = group_sum<L>(d_q);
            c2_q = group_sum<L>(c2_q);
            ss_q = group_sum<L>(ss_q);
            cnts = group_sum_u32<L>(cnts);
            const int n0 = cnts & 255u, n1 = (cnts >> 8) & 255u, n = n0 + n1;
            int status = TS_ST_OK;
            if ((cnts >> 16) != 0u || n0 == 0 || n1 == 0) status = TS_ST_VALUE_ERROR;
            const double margin = n == P ? margin_full : a.prm.draw_margin[n];   // ragged teams are rare: no indexed load on the hot path
            double v_s = 0.0, w_s = 0.0, v_q = 0.0, w_q = 0.0;
            const double rc_s = rsqrt(c2_s), rc_q = rsqrt(c2_q);
            if (status == TS_ST_OK &&
                (!trunc_vw(d_s * rc_s, margin * rc_s, draw_eff, a.prm.draw_is_zero, v_s, w_s) ||
                 !trunc_vw(d_q * rc_q, margin * rc_q, draw_eff, a.prm.draw_is_zero, v_q, w_q)))
                status = TS_ST_FP_ERROR;
            const bool ok = status == TS_ST_OK;
#pragma unroll
            for (int k = 0; k < SPL; ++k) {
                const int s = lane + k * L;
                const size_t o = (size_t)m * P + s;
                // a failed match writes nothing in the reference; here it republishes the
                // row it read under its own version so that successors can go on
                const double si = fma(es_sg[k], es_sg[k], tau2);
                const double gs = si * rc_s;
                const double ms2 = ok ? fma(sgn[k] * gs, v_s, es_mu[k]) : s_mu[k];
                const double ss2 = ok ? sqrt(si * (1.0 - (gs * rc_s) * w_s)) : s_sg[k];
                const double qi = fma(eq_sg[k], eq_sg[k], tau2);
                const double gq = qi * rc_q;
                const double mq2 = ok ? fma(sgn[k] * gq, v_q, eq_mu[k]) : q_mu[k];
                const double sq2 = ok ? sqrt(qi * (1.0 - (gq * rc_q) * w_q)) : q_sg[k];
                if (!(go && used[k])) continue;
                unsigned long long *row = a.vt + (size_t)pid[k] * row_words;
                st_row(row, ms2, ss2, a.seq_base + m);          // first: successors are waiting
                if (MULTI) {
                    // the queue pair moves on to its next ordinal; the ordinals travel with the shared pair
                    const bool upper = qj >= 3u;
                    uint32_t word = upper ? ord_hi[k] : ord_lo[k];
                    const uint32_t next = (((word >> ord_shift) & kOrdMask) + 1u) & kOrdMask;
                    word = (word & ~(kOrdMask << ord_shift)) | (next << ord_shift);
                    st_ord(row + 4, upper ? ord_lo[k] : word, upper ? word : ord_hi[k], a.seq_base + m);
                    st_row(row + q_word, mq2, sq2, next);
                } else {
                    st_row(row + 4, mq2, sq2, a.seq_base + m);
                }
                if (ok) {
                    if (a.out.out_shared) __stcg(reinterpret_cast<double2 *>(a.out.out_shared) + o, make_double2(ms2, ss2));
                    if (a.out.out_delta)   // rater.py:150-153
                        a.out.out_delta[o] = had[k] ? (ms2 - ss2) - (s_mu[k] - s_sg[k]) : 0.0;
                    if (a.out.out_queue) __stcg(reinterpret_cast<double2 *>(a.out.out_queue) + o, make_double2(mq2, sq2));
                }
            }
            if (go) {
                if (lane == 0) {
                    if (a.out.status) a.out.status[m] = (uint8_t)status;
                    if (!ok) atomicAdd(&a.ctrl[status == TS_ST_VALUE_ERROR ? CTRL_VALUE_ERRORS : CTRL_ERRORS], 1ull);
                    else if (a.out.quality) {      // rater.py:141: queue pairs before the update, no tau
                        const double nb2 = (double)n * beta2;
                        const double den = nb2 + ss_q;
                        a.out.quality[m] = exp(-0.5 * d_q * d_q / den) * sqrt(nb2 / den);
                    }
                }
                done = true;
            }
        }
    }
}

// user table rows [row_pairs][2] <-> versioned rows (shared column, queue column)
__global__ void __launch_bounds__(256)
k_vt_load_pair(const double *table, int row_pairs, int shared_col, int qcol, unsigned long long *vt, int64_t n)
{
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const double2 *row = reinterpret_cast<const double2 *>(table) + i * row_pairs;
    const double2 s = row[shared_col], q = row[qcol];
    st_row(vt + i * 8, s.x, s.y, TS_NONE);
    st_row(vt + i * 8 + 4, q.x, q.y, TS_NONE);
}
__global__ void __launch_bounds__(256)
k_vt_store_pair(const unsigned long long *vt, double *table, int row_pairs, int shared_col, int qcol, int64_t n)
{
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    unsigned long long ws[4], wq[4];
    ld_row(vt + i * 8, ws);
    ld_row(vt + i * 8 + 4, wq);
    double2 *row = reinterpret_cast<double2 *>(table) + i * row_pairs;
    row[shared_col] = make_double2(row_mu(ws), row_sigma(ws));
    row[qcol] = make_double2(row_mu(wq), row_sigma(wq));
}

// MULTI: table row [row_pairs][2] <-> shared pair, ordinals, the rated queue pairs
// (col_of_queue: 4 bits per compact queue index, its table column)
__global__ void __launch_bounds__(256)
k_vt_load_queues(const double *table, int row_pairs, int shared_col, uint32_t col_of_queue, int n_queues,
                 uint32_t row_words, unsigned long long *vt, int64_t n)
{
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const double2 *row = reinterpret_cast<const double2 *>(table) + i * row_pairs;
    unsigned long long *out = vt + i * row_words;
    const double2 sh = row[shared_col];
    st_row(out, sh.x, sh.y, TS_NONE);
    st_ord(out + 4, 0u, 0u, TS_NONE);
    for (int j = 0; j < n_queues; ++j) {
        const double2 q = row[(col_of_queue >> (4 * j)) & 15u];
        st_row(out + 8 + 4 * j, q.x, q.y, 0u);
    }
}
__global__ void __launch_bounds__(256)
k_vt_store_queues(const unsigned long long *vt, double *table, int row_pairs, int shared_col, uint32_t col_of_queue,
                  int n_queues, uint32_t row_words, int64_t n)
{
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    double2 *row = reinterpret_cast<double2 *>(table) + i * row_pairs;
    const unsigned long long *in = vt + i * row_words;
    unsigned long long w[4];
    ld_row(in, w);
    row[shared_col] = make_double2(row_mu(w), row_sigma(w));
    for (int j = 0; j < n_queues; ++j) {
        ld_row(in + 8 + 4 * j, w);
        row[(col_of_queue >> (4 * j)) & 15u] = make_double2(row_mu(w), row_sigma(w));
    }
}

template <int P, int L, bool MULTI>
int launch_pair(ts_ctx *ctx, const PairArgs &args, cudaStream_t stream)
{
    int per_sm = 0;
    TS_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_rate_stream_pair<P, L, MULTI>, kBlock, 0));
    if (per_sm < 1) { ts_set_error("stream kernel does not fit on an SM"); return TS_E_CUDA; }
    per_sm = per_sm < stream_blocks_per_sm_cap() ? per_sm : stream_blocks_per_sm_cap();
    dim3 grid(ctx->sm_count * per_sm), block(kBlock);
    void *kargs[] = {(void *)&args};
    TS_CUDA(cudaLaunchCooperativeKernel((void *)k_rate_stream_pair<P, L, MULTI>, grid, block, kargs, 0, stream));
    return 0;
}

template <int P, bool MULTI>
int dispatch_pair(ts_ctx *ctx, const PairArgs &a, int lanes, cudaStream_t s)
{
    switch (lanes) {
    case 1: return launch_pair<P, 1, MULTI>(ctx, a, s);
    case 2: return launch_pair<P, 2, MULTI>(ctx, a, s);
    case 4: return launch_pair<P, 4, MULTI>(ctx, a, s);
    case 8: return launch_pair<P, 8, MULTI>(ctx, a, s);
    case 16: return launch_pair<P, 16, MULTI>(ctx, a, s);
    case 32: return launch_pair<P, 32, MULTI>(ctx, a, s);
    }
    ts_set_error("lanes_per_match must be 0,1,2,4,8,16 or 32 (got %d)", lanes);
    return TS_E_ARG;
}

void fill_args(PairArgs &a, ts_ctx *ctx, const ts_table *table, const uint32_t *ids, const uint32_t *meta,
               const uint32_t *pred, int64_t n_matches, int64_t seq_base, const ts_outputs *outputs, void *workspace)
{
    a.prm = ctx->params;
    a.vt = (unsigned long long *)workspace;
    a.seed = table->seed;
    for (int i = 0; i < 16; ++i) a.modes.col[i] = i < 6 ? table->mode_col[i] : (int8_t)-1;
    a.ids = ids;
    a.meta = meta;
    a.pred = pred;
    a.row_words = 8;
    a.queue_nibbles = 0;
    a.n_matches = (uint32_t)n_matches;
    a.n_players = (uint32_t)(table->n_players > 0xFFFFFFFFll ? 0xFFFFFFFFll : table->n_players);
    a.seq_base = (uint32_t)seq_base;
    a.sleep_ns = stream_sleep_ns();
    a.max_spins = stream_max_spins();
    a.prefetch = stream_prefetch_mode() != 0 ? 1u : 0u;
    if (outputs) a.out = *outputs;
    else a.out = ts_outputs{nullptr, nullptr, nullptr, nullptr, nullptr};
    a.ctrl = ctx->d_ctrl;
}

// compact index over the distinct rated queue columns of a table (ascending column order)
struct QueueMap {
    int n_queues;
    uint32_t col_of_queue;    // 4 bits per compact index
    uint32_t queue_nibbles;   // 4 bits per game mode
};

int map_queues(const ts_table *table, QueueMap &qm)
{
    qm.n_queues = 0;
    qm.col_of_queue = qm.queue_nibbles = 0;
    int index_of_col[16];
    for (int c = 0; c < 16; ++c) index_of_col[c] = -1;
    for (int c = 0; c < 16; ++c) {
        bool rated = false;
        for (int i = 0; i < 6; ++i) rated |= table->mode_col[i] == c;
        if (!rated) continue;
        index_of_col[c] = qm.n_queues;
        qm.col_of_queue |= (uint32_t)c << (4 * qm.n_queues);
        ++qm.n_queues;
    }
    for (int i = 0; i < 6; ++i) {
        const int c = table->mode_col[i];
        if (c >= 0 && c < 16) qm.queue_nibbles |= (uint32_t)index_of_col[c] << (4 * i);
    }
    return qm.n_queues;
}

}  // namespace

// called by ts_rate_stream for tables with a shared column (stream.cu)
int ts_rate_stream_pair_impl(ts_ctx *ctx, const ts_table *table, int qcol, const uint32_t *ids, const uint32_t *meta,
                             const uint32_t *pred, int64_t n_matches, int32_t slots, int64_t seq_base, int32_t flags,
                             const ts_outputs *outputs, int32_t lanes_per_match, void *workspace, cudaStream_t s)
{
    PairArgs a;
    fill_args(a, ctx, table, ids, meta, pred, n_matches, seq_base, outputs, workspace);
    const int64_t u = table->n_players;
    if (flags & TS_STREAM_LOAD)
        k_vt_load_pair<<<(unsigned)((u + 255) / 256), 256, 0, s>>>(table->data, table->row_pairs, table->shared_col, qcol,
                                                                  a.vt, u);
    int lanes = lanes_per_match ? lanes_per_match : (slots == 6 ? 2 : 8);
    int rc = 0;
    if (n_matches > 0) rc = slots == 6 ? dispatch_pair<6, false>(ctx, a, lanes, s) : dispatch_pair<10, false>(ctx, a, lanes, s);
    if (rc) return rc;
    if (flags & TS_STREAM_STORE)
        k_vt_store_pair<<<(unsigned)((u + 255) / 256), 256, 0, s>>>(a.vt, table->data, table->row_pairs, table->shared_col,
                                                                   qcol, u);
    TS_CUDA(cudaGetLastError());
    return 0;
}

// versioned rows of a shared + several-queues table: shared pair, ordinals, one pair per rated queue column
int64_t ts_rate_stream_queues_row_bytes(const ts_table *table)
{
    QueueMap qm;
    return 8 * (8 + 4 * (int64_t)map_queues(table, qm));
}

// called by ts_rate_stream for tables with a shared column and several rated queue columns (stream.cu)
int ts_rate_stream_queues_impl(ts_ctx *ctx, const ts_table *table, const uint32_t *ids, const uint32_t *meta,
                               const uint32_t *pred, int64_t n_matches, int32_t slots, int64_t seq_base, int32_t flags,
                               const ts_outputs *outputs, int32_t lanes_per_match, void *workspace, cudaStream_t s)
{
    QueueMap qm;
    map_queues(table, qm);
    PairArgs a;
    fill_args(a, ctx, table, ids, meta, pred, n_matches, seq_base, outputs, workspace);
    a.row_words = 8u + 4u * (uint32_t)qm.n_queues;
    a.queue_nibbles = qm.queue_nibbles;
    const int64_t u = table->n_players;
    if ((flags & TS_STREAM_LOAD) && u > 0)
        k_vt_load_queues<<<(unsigned)((u + 255) / 256), 256, 0, s>>>(table->data, table->row_pairs, table->shared_col,
                                                                    qm.col_of_queue, qm.n_queues, a.row_words, a.vt, u);
    int lanes = lanes_per_match ? lanes_per_match : (slots == 6 ? 2 : 8);
    int rc = 0;
    if (n_matches > 0) rc = slots == 6 ? dispatch_pair<6, true>(ctx, a, lanes, s) : dispatch_pair<10, true>(ctx, a, lanes, s);
    if (rc) return rc;
    if ((flags & TS_STREAM_STORE) && u > 0)
        k_vt_store_queues<<<(unsigned)((u + 255) / 256), 256, 0, s>>>(a.vt, table->data, table->row_pairs, table->shared_col,
                                                                     qm.col_of_queue, qm.n_queues, a.row_words, u);
    TS_CUDA(cudaGetLastError());
    return 0;
}
