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
// MULTI (ts_rate_stream_queues): the batch rates matches of several game modes against
// a table with several queue columns (the faithful 8-pair row).  A player's shared
// pair is rewritten by every match of the player, a queue pair only by the matches
// of that queue, so the two carry different versions: the workspace mirrors the
// table (row_pairs x 4 words per player, every pair with its own version) and a
// slot waits for its shared pair at pred (previous match of the player) and for its
// queue pair at pred_queue (previous match of the player in the same queue column).
#include "stream_common.cuh"

namespace {

using namespace tsstream;
using tsvw::trunc_vw;

constexpr int kBlock = 256;

struct PairArgs {
    ts_params prm;
    unsigned long long *vt;       // 8 words per player
    const double *seed;
    ModeCols modes;               // >= 0: the mode is rated (all rated modes share the one queue column)
    const uint32_t *ids;
    const uint32_t *meta;
    const uint32_t *pred;
    const uint32_t *pred_q;       // MULTI only
    uint32_t row_words;           // words per player row: 8, MULTI: 4 * row_pairs
    uint32_t shared_word;         // word offset of the shared pair in a row: 0, MULTI: 4 * shared_col
    uint32_t col_nibbles;         // MULTI: 4 bits per game mode, its table column
    uint32_t n_matches;
    uint32_t seq_base;
    uint32_t sleep_ns;
    uint32_t max_spins;
    ts_outputs out;
    unsigned long long *ctrl;
};

template <int P, int W, bool MULTI>
struct WarpStage {
    uint32_t ids[2][W * P];
    uint32_t pred[2][W * P];
    uint32_t pred_q[2][MULTI ? W * P : 2];
    uint32_t meta[2][W];
};

template <int P, int W, bool MULTI>
__device__ __forceinline__ void stage_warp_tile(WarpStage<P, W, MULTI> &st, int b, const PairArgs &a, uint32_t tile, int wl)
{
    const uint32_t base = tile * W;
    const uint32_t cnt = min((uint32_t)W, a.n_matches - base);
    const uint32_t *src_i = a.ids + (size_t)base * P, *src_p = a.pred + (size_t)base * P;
    for (uint32_t i = wl; i < cnt * (P / 2); i += 32) {
        cp_async_8(&st.ids[b][2 * i], src_i + 2 * i);
        cp_async_8(&st.pred[b][2 * i], src_p + 2 * i);
        if (MULTI) cp_async_8(&st.pred_q[b][2 * i], a.pred_q + (size_t)base * P + 2 * i);
    }
    for (uint32_t i = wl; i < cnt; i += 32) cp_async_4(&st.meta[b][i], a.meta + base + i);
}

template <int P, int L, bool MULTI>
__global__ void __launch_bounds__(kBlock, 2) k_rate_stream_pair(const PairArgs a)
{
    constexpr int W = 32 / L;
    constexpr int SPL = (P + L - 1) / L;
    constexpr int HALF = P / 2;
    constexpr int kWarps = kBlock / 32;
    __shared__ __align__(16) WarpStage<P, W, MULTI> stage_all[kWarps];

    const int wl = threadIdx.x & 31;
    const int lane = wl % L;
    const int grp = wl / L;
    WarpStage<P, W, MULTI> &stage = stage_all[threadIdx.x >> 5];
    const uint32_t gmask = L == 32 ? 0xFFFFFFFFu : (((1u << L) - 1u) << (grp * L));
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

    int pf = 0;
    if (first_tile < n_tiles) stage_warp_tile<P, W, MULTI>(stage, pf, a, first_tile, wl);
    for (uint32_t tile = first_tile; tile < n_tiles; tile += stride) {
        cp_async_wait_all();
        __syncwarp();
        const int cur = pf;
        pf ^= 1;
        if (tile + stride < n_tiles) stage_warp_tile<P, W, MULTI>(stage, pf, a, tile + stride, wl);
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

        // word offsets of this match's two pairs in a player row
        const uint32_t row_words = MULTI ? a.row_words : 8u;
        const uint32_t s_word = MULTI ? a.shared_word : 0u;
        const uint32_t q_word = MULTI ? ((a.col_nibbles >> ((mode & 7u) * 4u)) & 7u) * 4u : 4u;
        uint32_t pid[SPL], want[SPL], want_q[SPL];
        bool used[SPL], have[SPL];
        double s_mu[SPL], s_sg[SPL], q_mu[SPL], q_sg[SPL];   // the row as read (NaN = NULL)
#pragma unroll
        for (int k = 0; k < SPL; ++k) {
            const int s = lane + k * L;
            pid[k] = (live && s < P) ? stage.ids[cur][grp * P + s] : TS_EMPTY_SLOT;
            used[k] = pid[k] != TS_EMPTY_SLOT;
            want[k] = used[k] ? stage.pred[cur][grp * P + s] : TS_NONE;
            want_q[k] = (MULTI && used[k]) ? stage.pred_q[cur][grp * P + s] : want[k];
            have[k] = !used[k];
            s_mu[k] = 0.0; s_sg[k] = 1.0; q_mu[k] = 0.0; q_sg[k] = 1.0;
        }
        bool done = !live;
        uint32_t spins = 0;
        while (!__all_sync(0xFFFFFFFFu, done)) {
            if ((++spins & 0x3FFu) == 0u) {      // watchdog, see stream.cu
                if (spins >= a.max_spins && wl == 0) atomicExch(&a.ctrl[CTRL_STALLED], 1ull);
                if (ld_relaxed_u64(&a.ctrl[CTRL_STALLED]) != 0ull) return;
            }
            if (!done) {
                bool ready = true;
#pragma unroll
                for (int k = 0; k < SPL; ++k) {
                    if (have[k]) continue;
                    unsigned long long ws[4], wq[4];
                    const unsigned long long *row = a.vt + (size_t)pid[k] * row_words;
                    ld_row(row + s_word, ws);
                    ld_row(row + q_word, wq);
                    if (row_has_version(ws, want[k]) && row_has_version(wq, want_q[k])) {
                        have[k] = true;
                        s_mu[k] = row_mu(ws); s_sg[k] = row_sigma(ws);
                        q_mu[k] = row_mu(wq); q_sg[k] = row_sigma(wq);
                    }
                    ready &= have[k];
                }
                if (__all_sync(gmask, ready)) {
                    double sgn[SPL], es_mu[SPL], es_sg[SPL], eq_mu[SPL], eq_sg[SPL];   // effective pairs
                    bool had[SPL];
                    double d_s = 0.0, c2_s = 0.0, d_q = 0.0, c2_q = 0.0, ss_q = 0.0;
                    uint32_t cnts = 0;
#pragma unroll
                    for (int k = 0; k < SPL; ++k) {
                        const int s = lane + k * L;
                        sgn[k] = ((s >= HALF ? 1 : 0) == first) ? 1.0 : -1.0;
                        es_mu[k] = s_mu[k]; es_sg[k] = s_sg[k]; eq_mu[k] = q_mu[k]; eq_sg[k] = q_sg[k];
                        had[k] = false;
                        if (!used[k]) continue;
                        bool bad = false;
                        had[k] = !isnan(s_mu[k]);
                        if (!had[k]) {                      // rater.py:118-119: first touch -> seed
                            double2 sd = make_double2(s_mu[k], s_sg[k]);
                            if (a.seed) sd = __ldg(reinterpret_cast<const double2 *>(a.seed) + pid[k]);
                            bad |= isnan(sd.x);
                            es_mu[k] = sd.x; es_sg[k] = sd.y;
                        }
                        if (isnan(q_mu[k])) { eq_mu[k] = es_mu[k]; eq_sg[k] = es_sg[k]; }   // rater.py:127-130
                        bad |= (eq_sg[k] == 0.0) || (es_sg[k] == 0.0);
                        d_s = fma(sgn[k], es_mu[k], d_s);
                        c2_s += fma(es_sg[k], es_sg[k], tb2);
                        d_q = fma(sgn[k], eq_mu[k], d_q);
                        c2_q += fma(eq_sg[k], eq_sg[k], tb2);
                        ss_q = fma(eq_sg[k], eq_sg[k], ss_q);
                        cnts += (sgn[k] > 0.0 ? 1u : 256u) + (bad ? 65536u : 0u);
                    }
                    if (L > 1) {
                        d_s = group_sum<L>(d_s, gmask);
                        c2_s = group_sum<L>(c2_s, gmask);
                        d_q = group_sum<L>(d_q, gmask);
                        c2_q = group_sum<L>(c2_q, gmask);
                        ss_q = group_sum<L>(ss_q, gmask);
                        cnts = group_sum_u32<L>(cnts, gmask);
                    }
                    const int n0 = cnts & 255u, n1 = (cnts >> 8) & 255u, n = n0 + n1;
                    int status = TS_ST_OK;
                    if ((cnts >> 16) != 0u || n0 == 0 || n1 == 0) status = TS_ST_VALUE_ERROR;
                    const double margin = n == P ? margin_full : a.prm.draw_margin[n];   // ragged teams are rare: no indexed load on the hot path
                    double v_s = 0.0, w_s = 0.0, v_q = 0.0, w_q = 0.0;
                    const double rc_s = rsqrt(c2_s), rc_q = rsqrt(c2_q);
                    if (status == TS_ST_OK &&
                        (!trunc_vw(d_s * rc_s, margin * rc_s, is_draw, a.prm.draw_is_zero != 0, v_s, w_s) ||
                         !trunc_vw(d_q * rc_q, margin * rc_q, is_draw, a.prm.draw_is_zero != 0, v_q, w_q)))
                        status = TS_ST_FP_ERROR;
#pragma unroll
                    for (int k = 0; k < SPL; ++k) {
                        if (!used[k]) continue;
                        const int s = lane + k * L;
                        const size_t o = (size_t)m * P + s;
                        // a failed match writes nothing in the reference; here it republishes the
                        // row it read under its own version so that successors can go on
                        double ms2 = s_mu[k], ss2 = s_sg[k], mq2 = q_mu[k], sq2 = q_sg[k];
                        if (status == TS_ST_OK) {
                            const double si = fma(es_sg[k], es_sg[k], tau2);
                            const double gs = si * rc_s;
                            ms2 = fma(sgn[k] * gs, v_s, es_mu[k]);
                            ss2 = sqrt(si * (1.0 - (gs * rc_s) * w_s));
                            const double qi = fma(eq_sg[k], eq_sg[k], tau2);
                            const double gq = qi * rc_q;
                            mq2 = fma(sgn[k] * gq, v_q, eq_mu[k]);
                            sq2 = sqrt(qi * (1.0 - (gq * rc_q) * w_q));
                        }
                        unsigned long long *row = a.vt + (size_t)pid[k] * row_words;
                        st_row(row + s_word, ms2, ss2, a.seq_base + m);          // first: successors are waiting
                        st_row(row + q_word, mq2, sq2, a.seq_base + m);
                        if (status == TS_ST_OK) {
                            if (a.out.out_shared) __stcg(reinterpret_cast<double2 *>(a.out.out_shared) + o, make_double2(ms2, ss2));
                            if (a.out.out_delta)   // rater.py:150-153
                                a.out.out_delta[o] = had[k] ? (ms2 - ss2) - (s_mu[k] - s_sg[k]) : 0.0;
                            if (a.out.out_queue) __stcg(reinterpret_cast<double2 *>(a.out.out_queue) + o, make_double2(mq2, sq2));
                        }
                    }
                    if (lane == 0) {
                        if (a.out.status) a.out.status[m] = (uint8_t)status;
                        if (status != TS_ST_OK) atomicAdd(&a.ctrl[CTRL_ERRORS], 1ull);
                        else if (a.out.quality) {      // rater.py:141: queue pairs before the update, no tau
                            const double nb2 = (double)n * beta2;
                            const double den = nb2 + ss_q;
                            a.out.quality[m] = exp(-0.5 * d_q * d_q / den) * sqrt(nb2 / den);
                        }
                    }
                    done = true;
                } else {
                    __nanosleep(a.sleep_ns);
                }
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

// MULTI: every pair of the row, each with its own version (one thread per pair)
__global__ void __launch_bounds__(256) k_vt_load_all(const double *table, unsigned long long *vt, int64_t n_pairs)
{
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_pairs) return;
    const double2 r = reinterpret_cast<const double2 *>(table)[i];
    st_row(vt + i * 4, r.x, r.y, TS_NONE);
}
__global__ void __launch_bounds__(256) k_vt_store_all(const unsigned long long *vt, double *table, int64_t n_pairs)
{
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_pairs) return;
    unsigned long long w[4];
    ld_row(vt + i * 4, w);
    reinterpret_cast<double2 *>(table)[i] = make_double2(row_mu(w), row_sigma(w));
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
    // one lane per match stages a 32-match tile: with the second link array that is past the
    // 48 KB of static shared memory for 5v5, and one lane is never the fast choice
    if constexpr (!MULTI) {
        if (lanes == 1) return launch_pair<P, 1, false>(ctx, a, s);
    }
    switch (lanes) {
    case 2: return launch_pair<P, 2, MULTI>(ctx, a, s);
    case 4: return launch_pair<P, 4, MULTI>(ctx, a, s);
    case 8: return launch_pair<P, 8, MULTI>(ctx, a, s);
    case 16: return launch_pair<P, 16, MULTI>(ctx, a, s);
    case 32: return launch_pair<P, 32, MULTI>(ctx, a, s);
    }
    ts_set_error(MULTI ? "lanes_per_match must be 0,2,4,8,16 or 32 (got %d)" : "lanes_per_match must be 0,1,2,4,8,16 or 32 (got %d)", lanes);
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
    a.pred_q = nullptr;
    a.row_words = 8;
    a.shared_word = 0;
    a.col_nibbles = 0;
    a.n_matches = (uint32_t)n_matches;
    a.seq_base = (uint32_t)seq_base;
    a.sleep_ns = stream_sleep_ns();
    a.max_spins = stream_max_spins();
    if (outputs) a.out = *outputs;
    else a.out = ts_outputs{nullptr, nullptr, nullptr, nullptr, nullptr};
    a.ctrl = ctx->d_ctrl;
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
    TS_CUDA(cudaMemsetAsync(ctx->d_ctrl + CTRL_STALLED, 0, sizeof(unsigned long long), s));
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

extern "C" int64_t ts_rate_stream_queues_workspace_bytes(int64_t n_players, int32_t row_pairs)
{
    if (n_players < 0 || row_pairs < 2 || row_pairs > 8) {
        ts_set_error("ts_rate_stream_queues: the table needs a shared column and 2..8 pairs per row");
        return TS_E_ARG;
    }
    return n_players * row_pairs * 32;
}

extern "C" int ts_rate_stream_queues(ts_ctx *ctx, const ts_table *table, const uint32_t *ids, const uint32_t *meta,
                                     const uint32_t *pred, const uint32_t *pred_queue, int64_t n_matches, int32_t slots,
                                     int64_t seq_base, int32_t flags, const ts_outputs *outputs, int32_t lanes_per_match,
                                     void *workspace, int64_t workspace_bytes, void *stream)
{
    if (seq_base < 0 || n_matches < 0 || seq_base + n_matches >= 0xFFFFFFFFll) {
        ts_set_error("ts_rate_stream_queues: stream-wide match numbers must stay below 2^32 - 1");
        return TS_E_ARG;
    }
    if (!ctx || !table || !table->data || (n_matches > 0 && (!ids || !meta || !pred || !pred_queue || !workspace))) {
        ts_set_error("ts_rate_stream_queues: null argument");
        return TS_E_ARG;
    }
    if (table->row_pairs < 2 || table->row_pairs > 8 || table->shared_col < 0 || table->shared_col >= table->row_pairs) {
        ts_set_error("ts_rate_stream_queues: the table needs a shared column and 2..8 pairs per row");
        return TS_E_UNSUPPORTED;
    }
    uint32_t nibbles = 0;
    for (int i = 0; i < 6; ++i) {
        const int c = table->mode_col[i];
        if (c < 0) continue;
        if (c >= table->row_pairs || c == table->shared_col) { ts_set_error("ts_rate_stream_queues: bad mode_col[%d]", i); return TS_E_ARG; }
        nibbles |= (uint32_t)c << (4 * i);
    }
    if (slots != 6 && slots != 10) {
        ts_set_error("ts_rate_stream_queues: slots must be 6 or 10 (got %d)", slots);
        return TS_E_UNSUPPORTED;
    }
    const int64_t need = ts_rate_stream_queues_workspace_bytes(table->n_players, table->row_pairs);
    if (workspace_bytes < need) {
        ts_set_error("ts_rate_stream_queues: workspace too small (%lld < %lld bytes)", (long long)workspace_bytes, (long long)need);
        return TS_E_WORKSPACE;
    }
    TS_CUDA(cudaSetDevice(ctx->device));
    cudaStream_t s = (cudaStream_t)stream;
    PairArgs a;
    fill_args(a, ctx, table, ids, meta, pred, n_matches, seq_base, outputs, workspace);
    a.pred_q = pred_queue;
    a.row_words = 4u * (uint32_t)table->row_pairs;
    a.shared_word = 4u * (uint32_t)table->shared_col;
    a.col_nibbles = nibbles;
    const int64_t pairs = table->n_players * table->row_pairs;
    if ((flags & TS_STREAM_LOAD) && pairs > 0)
        k_vt_load_all<<<(unsigned)((pairs + 255) / 256), 256, 0, s>>>(table->data, a.vt, pairs);
    TS_CUDA(cudaMemsetAsync(ctx->d_ctrl + CTRL_STALLED, 0, sizeof(unsigned long long), s));
    int lanes = lanes_per_match ? lanes_per_match : (slots == 6 ? 2 : 8);
    int rc = 0;
    if (n_matches > 0) rc = slots == 6 ? dispatch_pair<6, true>(ctx, a, lanes, s) : dispatch_pair<10, true>(ctx, a, lanes, s);
    if (rc) return rc;
    if ((flags & TS_STREAM_STORE) && pairs > 0)
        k_vt_store_all<<<(unsigned)((pairs + 255) / 256), 256, 0, s>>>(a.vt, table->data, pairs);
    TS_CUDA(cudaGetLastError());
    return 0;
}
