// Streaming (dataflow) form of the fp64 rating kernel for sm_100a.
//
// Same arithmetic and same result as ts_rate, but no rounds and no grid
// barriers: matches are rated in stream order (worker.py:192's order, which is
// a topological order of the dependency graph), and a match simply waits until
// the rows of its players carry the version it expects -- the index of that
// player's previous match, from ts_link_pred -- before it reads them.  On B200
// a coherent L2 round trip costs ~0.6 us, so a barrier per round puts several
// of them on the critical path of each of the R rounds (R = 888 for 10M
// matches over 1M players); here the only chained latency is one poll per hop
// of the longest dependency chain.
//
// Row format (workspace, 32 bytes = one sector per player): four 64-bit words
//   [mu.lo | ver<<32] [mu.hi | ver<<32] [sigma.lo | ver<<32] [sigma.hi | ver<<32]
// Every 8-byte word validates itself (64-bit accesses are single-copy atomic),
// so neither writer nor reader needs a fence: a reader accepts a row when all
// four version fields equal the version it waits for.  Same idea as the
// low-latency flag protocols of collective libraries.
//
// Progress: warps take tiles of consecutive matches round-robin in increasing
// order and the grid is co-resident (cooperative launch), so the oldest
// unfinished match never waits on unstarted work.  Inside a warp the wait loop
// is warp-uniform and complete groups finish whole matches inside an iteration,
// so a match may depend on another match of the same warp.
//
// This file: single-pair ladders (table.row_pairs == 1, shared_col == -1) and the
// ts_rate_stream entry point; the shared + queue layouts are in stream_pair.cu.
#include "stream_common.cuh"

namespace {

using namespace tsstream;

#ifndef TS_STREAM_BLOCK
#define TS_STREAM_BLOCK 256
#endif
constexpr int kBlock = TS_STREAM_BLOCK;
#ifndef TS_STREAM_MIN_BLOCKS
#define TS_STREAM_MIN_BLOCKS 2   // <= 128 registers: 2 blocks/SM; measured config 2 rate: 1 block 3.02, 2 blocks 2.19, 3 blocks (80 registers) 2.30 ms
#endif

constexpr int kMaxShards = 16;

struct StreamArgs {
    ts_params prm;
    unsigned long long *vt;       // versioned rows, 4 words per player (single-GPU form)
    // sharded form: player p lives on GPU p % n_shards, local row p / n_shards; shards[g] is
    // that GPU's row array (own memory or a peer mapping reached over NVLink)
    unsigned long long *shards[kMaxShards];
    uint32_t n_shards;            // power of two
    uint32_t shard_shift;         // log2(n_shards)
    uint32_t rank;
    const double *seed;
    ModeCols modes;
    const uint32_t *ids;
    const uint32_t *meta;
    const uint32_t *pred;
    uint32_t n_matches;
    uint32_t seq_base;            // stream-wide number of the first match of this window
    uint32_t sleep_ns;            // back-off between polls when no group of the warp is complete
    uint32_t max_spins;           // watchdog: fruitless polls before a warp declares the stream stalled
    ts_outputs out;
    unsigned long long *ctrl;
};

template <bool SHARDED>
__device__ __forceinline__ unsigned long long *row_of(const StreamArgs &a, uint32_t pid)
{
    if (!SHARDED) return a.vt + (size_t)pid * 4;
    return a.shards[pid & (a.n_shards - 1u)] + (size_t)(pid >> a.shard_shift) * 4;
}

using tsvw::trunc_vw;

// 1: the loads of all missing slots of a poll are issued before any is checked.  On one GPU
// (rows in the local L2) checking slot by slot measured slightly faster (config 2 rate 2.19 vs
// 2.25 ms); rows in peer memory are a NVLink round trip away and are always fetched together.
#ifndef TS_STREAM_HOIST
#define TS_STREAM_HOIST 0
#endif

// per-warp staging buffers: a warp owns W = 32 / L consecutive matches at a time
template <int P, int W>
struct WarpStage {
    uint32_t ids[2][W * P];
    uint32_t pred[2][W * P];
    uint32_t meta[2][W];
};

template <int P, int W>
__device__ __forceinline__ void stage_warp_tile(WarpStage<P, W> &st, int b, const StreamArgs &a, uint32_t tile, int wl)
{
    const uint32_t base = tile * W;
    const uint32_t cnt = min((uint32_t)W, a.n_matches - base);
    const uint32_t *src_i = a.ids + (size_t)base * P, *src_p = a.pred + (size_t)base * P;
    for (uint32_t i = wl; i < cnt * (P / 2); i += 32) {
        cp_async_8(&st.ids[b][2 * i], src_i + 2 * i);
        cp_async_8(&st.pred[b][2 * i], src_p + 2 * i);
    }
    for (uint32_t i = wl; i < cnt; i += 32) cp_async_4(&st.meta[b][i], a.meta + base + i);
}

// Warps are the unit of work: warp w of the grid takes the tiles w, w + n_warps,
// ... of W consecutive matches, in increasing order, and never synchronises
// with the other warps of its block -- a warp that has to wait for a dependency
// holds up nothing but its own W matches.
// SHARDED: this GPU rates the tiles t with t % n_shards == rank of ONE ladder whose
// rows are spread over all GPUs; the version protocol is the only synchronisation
// between GPUs (no inter-GPU barrier, no collective).
template <int P, int L, bool SHARDED>
__global__ void __launch_bounds__(kBlock, TS_STREAM_MIN_BLOCKS) k_rate_stream(const StreamArgs a)
{
    constexpr int W = 32 / L;
    constexpr int SPL = (P + L - 1) / L;
    constexpr int HALF = P / 2;
    constexpr int kWarps = kBlock / 32;
    __shared__ __align__(16) WarpStage<P, W> stage_all[kWarps];

    const int wl = threadIdx.x & 31;            // lane in warp
    const int lane = wl % L;                    // lane in match group
    const int grp = wl / L;                     // match of the warp tile
    WarpStage<P, W> &stage = stage_all[threadIdx.x >> 5];
    const double tau2 = a.prm.tau * a.prm.tau;
    const double beta2 = a.prm.beta * a.prm.beta;
    const double tb2 = tau2 + beta2;
    const double margin_full = a.prm.draw_margin[P];
    uint32_t rated_mask = 0;
#pragma unroll
    for (int i = 0; i < 6; ++i) rated_mask |= a.modes.col[i] >= 0 ? (1u << i) : 0u;
    const uint32_t n_tiles = (a.n_matches + W - 1) / W;
    const uint32_t stride = gridDim.x * kWarps * (SHARDED ? a.n_shards : 1u);
    // neighbouring tiles go to different GPUs, then to different SMs
    const uint32_t first = ((threadIdx.x >> 5) * gridDim.x + blockIdx.x) * (SHARDED ? a.n_shards : 1u) + (SHARDED ? a.rank : 0u);

    int pf = 0;
    if (first < n_tiles) stage_warp_tile<P, W>(stage, pf, a, first, wl);
    for (uint32_t tile = first; tile < n_tiles; tile += stride) {
        cp_async_wait_all();
        __syncwarp();
        const int cur = pf;
        pf ^= 1;
        if (tile + stride < n_tiles) stage_warp_tile<P, W>(stage, pf, a, tile + stride, wl);
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

        uint32_t pid[SPL], want[SPL];
        bool used[SPL], have[SPL];
        double q_mu[SPL], q_sg[SPL];
#pragma unroll
        for (int k = 0; k < SPL; ++k) {
            const int s = lane + k * L;
            pid[k] = (live && s < P) ? stage.ids[cur][grp * P + s] : TS_EMPTY_SLOT;
            used[k] = pid[k] != TS_EMPTY_SLOT;
            want[k] = used[k] ? stage.pred[cur][grp * P + s] : TS_NONE;
            have[k] = !used[k];
            q_mu[k] = 0.0; q_sg[k] = 1.0;
        }
        bool done = !live;
        // Wait loop, the same for all lanes of the warp.  Each iteration polls the rows still
        // missing; then ONE ballot tells every lane which match groups are complete, and if
        // there is any, the whole warp runs the rating arithmetic together -- the issue slots
        // are spent whether 2 or 32 lanes are active, and a converged warp needs no
        // reconvergence barriers, no partial-mask votes and no mask checks on the shuffles;
        // only the stores are predicated on the lane's group being complete.  Lanes of
        // groups that are not complete compute on placeholders and throw the result away.
        // Watchdog: links that can never be satisfied (a pred array that does not belong to
        // this stream) must not hang the GPU -- after max_spins (2^24) fruitless polls (tens of
        // seconds; a legitimate wait is milliseconds) a warp raises the stall flag, and every
        // warp that sees it gives up.
        uint32_t spins = 0;
        while (!__all_sync(0xFFFFFFFFu, done)) {
            bool ready = false;
            if (!done) {
                ready = true;
                if constexpr (SHARDED || TS_STREAM_HOIST) {
                    unsigned long long w[SPL][4];
#pragma unroll
                    for (int k = 0; k < SPL; ++k)
                        if (!have[k]) ld_row<SHARDED>(row_of<SHARDED>(a, pid[k]), w[k]);
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
                        ld_row<SHARDED>(row_of<SHARDED>(a, pid[k]), w);
                        if (row_has_version(w, want[k])) {
                            have[k] = true;
                            q_mu[k] = row_mu(w);
                            q_sg[k] = row_sigma(w);
                        }
                        ready &= have[k];
                    }
                }
            }
            const uint32_t complete = complete_groups<L>(ready);
            if (complete == 0u) {
                if ((++spins & 0x3FFu) == 0u) {
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

            double sgn[SPL], e_mu[SPL], e_sg[SPL];     // effective pairs; q_mu / q_sg stay the rows as read
            double d_q = 0.0, c2_q = 0.0, ss_q = 0.0;
            uint32_t cnts = 0;
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
                    if (a.seed) sd = __ldg(reinterpret_cast<const double2 *>(a.seed) + pid[k]);
                    if (isnan(sd.x)) cnts |= 65536u;       // neither rating nor seed
                    else { e_mu[k] = sd.x; e_sg[k] = sd.y; }
                }
            }
#pragma unroll
            for (int k = 0; k < SPL; ++k) {
                const int s = lane + k * L;
                sgn[k] = ((s >= HALF ? 1 : 0) == first) ? 1.0 : -1.0;
                const bool u = used[k];
                d_q = fma(sgn[k], u ? e_mu[k] : 0.0, d_q);
                c2_q += u ? fma(e_sg[k], e_sg[k], tb2) : 0.0;
                ss_q = u ? fma(e_sg[k], e_sg[k], ss_q) : ss_q;
                cnts += (u ? (sgn[k] > 0.0 ? 1u : 256u) : 0u) + ((u && e_sg[k] == 0.0) ? 65536u : 0u);
            }
            d_q = group_sum<L>(d_q);
            c2_q = group_sum<L>(c2_q);
            ss_q = group_sum<L>(ss_q);
            cnts = group_sum_u32<L>(cnts);
            const int n0 = cnts & 255u, n1 = (cnts >> 8) & 255u, n = n0 + n1;
            int status = TS_ST_OK;
            if ((cnts >> 16) != 0u || n0 == 0 || n1 == 0) status = TS_ST_VALUE_ERROR;
            const double margin = n == P ? margin_full : a.prm.draw_margin[n];   // ragged teams are rare: no indexed load on the hot path
            double v_q = 0.0, w_q = 0.0;
            const double rc_q = rsqrt(c2_q);
            if (status == TS_ST_OK &&
                !trunc_vw(d_q * rc_q, margin * rc_q, draw_eff, a.prm.draw_is_zero != 0, v_q, w_q))
                status = TS_ST_FP_ERROR;
            const bool ok = status == TS_ST_OK;
#pragma unroll
            for (int k = 0; k < SPL; ++k) {
                const int s = lane + k * L;
                // a failed match writes nothing in the reference; here it republishes the
                // row it read under its own version so that successors can go on
                const double si = fma(e_sg[k], e_sg[k], tau2);
                const double g = si * rc_q;
                const double mu2 = ok ? fma(sgn[k] * g, v_q, e_mu[k]) : q_mu[k];
                const double sg2 = ok ? sqrt(si * (1.0 - (g * rc_q) * w_q)) : q_sg[k];
                if (go && used[k]) {
                    st_row<SHARDED>(row_of<SHARDED>(a, pid[k]), mu2, sg2, a.seq_base + m);   // first: successors are waiting
                    if (ok && a.out.out_queue)
                        __stcg(reinterpret_cast<double2 *>(a.out.out_queue) + (size_t)m * P + s, make_double2(mu2, sg2));
                }
            }
            if (go) {
                if (lane == 0) {
                    if (a.out.status) a.out.status[m] = (uint8_t)status;
                    if (!ok) atomicAdd(&a.ctrl[CTRL_ERRORS], 1ull);
                    else if (a.out.quality) {
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

// user table (mu, sigma) <-> versioned rows
__global__ void __launch_bounds__(256) k_vt_load(const double *table, unsigned long long *vt, int64_t n)
{
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    double2 r = reinterpret_cast<const double2 *>(table)[i];
    st_row(vt + i * 4, r.x, r.y, TS_NONE);
}
__global__ void __launch_bounds__(256) k_vt_store(const unsigned long long *vt, double *table, int64_t n)
{
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    unsigned long long w[4];
    ld_row(vt + i * 4, w);
    reinterpret_cast<double2 *>(table)[i] = make_double2(row_mu(w), row_sigma(w));
}

template <int P, int L, bool SHARDED>
int launch_stream(ts_ctx *ctx, const StreamArgs &args, cudaStream_t stream)
{
    int per_sm = 0;
    TS_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_rate_stream<P, L, SHARDED>, kBlock, 0));
    if (per_sm < 1) { ts_set_error("stream kernel does not fit on an SM"); return TS_E_CUDA; }
    per_sm = per_sm < stream_blocks_per_sm_cap() ? per_sm : stream_blocks_per_sm_cap();
    dim3 grid(ctx->sm_count * per_sm), block(kBlock);
    void *kargs[] = {(void *)&args};
    // cooperative launch only for its guarantee that the whole grid is resident
    TS_CUDA(cudaLaunchCooperativeKernel((void *)k_rate_stream<P, L, SHARDED>, grid, block, kargs, 0, stream));
    return 0;
}

template <int P, bool SHARDED>
int dispatch_l(ts_ctx *ctx, const StreamArgs &a, int lanes, cudaStream_t s)
{
    switch (lanes) {
    case 1: return launch_stream<P, 1, SHARDED>(ctx, a, s);
    case 2: return launch_stream<P, 2, SHARDED>(ctx, a, s);
    case 4: return launch_stream<P, 4, SHARDED>(ctx, a, s);
    case 8: return launch_stream<P, 8, SHARDED>(ctx, a, s);
    case 16: return launch_stream<P, 16, SHARDED>(ctx, a, s);
    case 32: return launch_stream<P, 32, SHARDED>(ctx, a, s);
    }
    ts_set_error("lanes_per_match must be 0,1,2,4,8,16 or 32 (got %d)", lanes);
    return TS_E_ARG;
}

}  // namespace

// 64 bytes per player covers two row formats (one pair: 32, shared + one queue pair: 64)
extern "C" int64_t ts_rate_stream_workspace_bytes(int64_t n_players) { return n_players < 0 ? TS_E_ARG : n_players * 64; }

int64_t ts_rate_stream_queues_row_bytes(const ts_table *table);
int ts_rate_stream_queues_impl(ts_ctx *ctx, const ts_table *table, const uint32_t *ids, const uint32_t *meta,
                               const uint32_t *pred, int64_t n_matches, int32_t slots, int64_t seq_base, int32_t flags,
                               const ts_outputs *outputs, int32_t lanes_per_match, void *workspace, cudaStream_t s);

namespace {
// rated queue columns of a table: how many distinct ones, and the last one seen
int rated_queue_columns(const ts_table *table, int *last_col)
{
    uint32_t seen = 0;
    int n = 0;
    for (int i = 0; i < 6; ++i) {
        const int c = table->mode_col[i];
        if (c < 0 || c > 15) continue;
        if (!((seen >> c) & 1u)) { seen |= 1u << c; ++n; }
        if (last_col) *last_col = c;
    }
    return n;
}
}  // namespace

// any layout ts_rate_stream takes: 32 (one pair), 64 (shared + one queue) or 64 + 32 per rated
// queue column (shared + several queues) bytes per player
extern "C" int64_t ts_rate_stream_table_workspace_bytes(const ts_table *table)
{
    if (!table || table->n_players < 0) return TS_E_ARG;
    if (table->shared_col < 0) return table->n_players * 32;
    if (rated_queue_columns(table, nullptr) <= 1) return table->n_players * 64;
    return table->n_players * ts_rate_stream_queues_row_bytes(table);
}

int ts_rate_stream_pair_impl(ts_ctx *ctx, const ts_table *table, int qcol, const uint32_t *ids, const uint32_t *meta,
                             const uint32_t *pred, int64_t n_matches, int32_t slots, int64_t seq_base, int32_t flags,
                             const ts_outputs *outputs, int32_t lanes_per_match, void *workspace, cudaStream_t s);

extern "C" int ts_stream_stalled(ts_ctx *ctx, void *stream)
{
    if (!ctx) return TS_E_ARG;
    unsigned long long v = 0;
    TS_CUDA(cudaMemcpyAsync(&v, ctx->d_ctrl + CTRL_STALLED, sizeof v, cudaMemcpyDeviceToHost, (cudaStream_t)stream));
    TS_CUDA(cudaStreamSynchronize((cudaStream_t)stream));
    return v != 0ull ? 1 : 0;
}

extern "C" int ts_rate_stream(ts_ctx *ctx, const ts_table *table, const uint32_t *ids, const uint32_t *meta,
                              const uint32_t *pred, int64_t n_matches, int32_t slots, int64_t seq_base, int32_t flags,
                              const ts_outputs *outputs, int32_t lanes_per_match, void *workspace,
                              int64_t workspace_bytes, void *stream)
{
    if (seq_base < 0 || seq_base + n_matches >= 0xFFFFFFFFll) {
        ts_set_error("ts_rate_stream: stream-wide match numbers must stay below 2^32 - 1");
        return TS_E_ARG;
    }
    if (!ctx || !table || !table->data || (n_matches > 0 && (!ids || !meta || !pred || !workspace))) {
        ts_set_error("ts_rate_stream: null argument");
        return TS_E_ARG;
    }
    // layouts: one pair per player (row_pairs == 1, shared_col == -1), or a shared pair plus queue columns
    for (int i = 0; i < 6; ++i) {
        const int c = table->mode_col[i];
        if (c >= 0 && (c >= table->row_pairs || c > 15 || c == table->shared_col)) {
            ts_set_error("ts_rate_stream: bad mode_col[%d]", i);
            return TS_E_ARG;
        }
    }
    int qcol = -1;
    const int n_queues = rated_queue_columns(table, &qcol);
    const bool pair = table->shared_col >= 0;
    if (pair ? (n_queues < 1 || table->shared_col >= table->row_pairs) : (table->row_pairs != 1)) {
        ts_set_error("ts_rate_stream: the table must be a single-pair ladder (row_pairs == 1, shared_col == -1) or a "
                     "shared pair plus at least one rated queue column");
        return TS_E_UNSUPPORTED;
    }
    if (slots != 6 && slots != 10) {
        ts_set_error("ts_rate_stream: slots must be 6 or 10 (got %d)", slots);
        return TS_E_UNSUPPORTED;
    }
    if (n_matches < 0 || n_matches >= 0xFFFFFFFFll) {
        ts_set_error("ts_rate_stream: n_matches outside [0, 2^32 - 1)");
        return TS_E_ARG;
    }
    if (workspace_bytes < ts_rate_stream_table_workspace_bytes(table)) {
        ts_set_error("ts_rate_stream: workspace too small (%lld < %lld bytes)", (long long)workspace_bytes,
                     (long long)ts_rate_stream_table_workspace_bytes(table));
        return TS_E_WORKSPACE;
    }
    TS_CUDA(cudaSetDevice(ctx->device));
    cudaStream_t s = (cudaStream_t)stream;
    if (pair && n_queues > 1)
        return ts_rate_stream_queues_impl(ctx, table, ids, meta, pred, n_matches, slots, seq_base, flags, outputs,
                                          lanes_per_match, workspace, s);
    if (pair)
        return ts_rate_stream_pair_impl(ctx, table, qcol, ids, meta, pred, n_matches, slots, seq_base, flags, outputs,
                                        lanes_per_match, workspace, s);
    StreamArgs a;
    a.prm = ctx->params;
    a.vt = (unsigned long long *)workspace;
    a.seed = table->seed;
    for (int i = 0; i < 16; ++i) a.modes.col[i] = i < 6 ? table->mode_col[i] : (int8_t)-1;
    a.ids = ids;
    a.meta = meta;
    a.pred = pred;
    a.n_matches = (uint32_t)n_matches;
    a.seq_base = (uint32_t)seq_base;
    a.sleep_ns = stream_sleep_ns();
    a.max_spins = stream_max_spins();
    if (outputs) a.out = *outputs;
    else a.out = ts_outputs{nullptr, nullptr, nullptr, nullptr, nullptr};
    a.ctrl = ctx->d_ctrl;
    const int64_t u = table->n_players;
    if (flags & TS_STREAM_LOAD) k_vt_load<<<(unsigned)((u + 255) / 256), 256, 0, s>>>(table->data, a.vt, u);
    int lanes = lanes_per_match ? lanes_per_match : (slots == 6 ? 2 : 8);   // measured: DESIGN.md 3.2
    a.n_shards = 1; a.shard_shift = 0; a.rank = 0;
    for (int i = 0; i < kMaxShards; ++i) a.shards[i] = nullptr;
    TS_CUDA(cudaMemsetAsync(ctx->d_ctrl + CTRL_STALLED, 0, sizeof(unsigned long long), s));
    int rc = 0;
    if (n_matches > 0) rc = slots == 6 ? dispatch_l<6, false>(ctx, a, lanes, s) : dispatch_l<10, false>(ctx, a, lanes, s);
    if (rc) return rc;
    if (flags & TS_STREAM_STORE) k_vt_store<<<(unsigned)((u + 255) / 256), 256, 0, s>>>(a.vt, table->data, u);
    TS_CUDA(cudaGetLastError());
    return 0;
}


// ---------------------------------------------------------------------------
// One ladder, table hash-sharded over the GPUs of a node (BASELINE.json
// configs[4]): player p is owned by GPU p % G (G a power of two), local row
// p / G.  Every GPU maps all shards (its own allocation plus its peers' through
// CUDA IPC, reached over NVLink) and rates the warp tiles t with t % G == rank.
// The row versions are the only synchronisation between GPUs.
// ---------------------------------------------------------------------------
namespace {

__global__ void __launch_bounds__(256)
k_shard_load(const double *table_shard, unsigned long long *vt_shard, int64_t rows)
{
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= rows) return;
    double2 r = reinterpret_cast<const double2 *>(table_shard)[i];
    st_row<true>(vt_shard + i * 4, r.x, r.y, TS_NONE);
}
__global__ void __launch_bounds__(256)
k_shard_store(const unsigned long long *vt_shard, double *table_shard, int64_t rows)
{
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= rows) return;
    unsigned long long w[4];
    ld_row<true>(vt_shard + i * 4, w);
    reinterpret_cast<double2 *>(table_shard)[i] = make_double2(row_mu(w), row_sigma(w));
}

}  // namespace

extern "C" int ts_shard_alloc(int device, int64_t bytes, void **ptr, void *handle64)
{
    if (!ptr || !handle64 || bytes <= 0) { ts_set_error("ts_shard_alloc: bad argument"); return TS_E_ARG; }
    static_assert(sizeof(cudaIpcMemHandle_t) == 64, "IPC handle size");
    TS_CUDA(cudaSetDevice(device));
    TS_CUDA(cudaMalloc(ptr, (size_t)bytes));
    TS_CUDA(cudaIpcGetMemHandle((cudaIpcMemHandle_t *)handle64, *ptr));
    return 0;
}
extern "C" int ts_shard_open(int device, const void *handle64, void **ptr)
{
    if (!ptr || !handle64) { ts_set_error("ts_shard_open: bad argument"); return TS_E_ARG; }
    cudaIpcMemHandle_t h;
    memcpy(&h, handle64, sizeof h);
    TS_CUDA(cudaSetDevice(device));
    TS_CUDA(cudaIpcOpenMemHandle(ptr, h, cudaIpcMemLazyEnablePeerAccess));
    return 0;
}
extern "C" int ts_shard_close(void *ptr)
{
    if (ptr) TS_CUDA(cudaIpcCloseMemHandle(ptr));
    return 0;
}
extern "C" int ts_shard_free(void *ptr)
{
    if (ptr) TS_CUDA(cudaFree(ptr));
    return 0;
}

// phase 0: table_shard -> own rows; phase 1: rate this rank's tiles; phase 2: own rows -> table_shard.
// The caller separates the phases with a barrier across the ranks (every shard must be loaded
// before any rank polls it, and every rank must have finished before rows are read back).
extern "C" int ts_rate_stream_sharded(ts_ctx *ctx, int32_t phase, double *table_shard, int64_t n_players,
                                      int32_t rank, int32_t n_shards, void *const *vt_shards, int8_t mode,
                                      const uint32_t *ids, const uint32_t *meta, const uint32_t *pred,
                                      int64_t n_matches, int32_t slots, const ts_outputs *outputs,
                                      int32_t lanes_per_match, void *stream)
{
    if (!ctx || !vt_shards || n_shards < 1 || n_shards > kMaxShards || (n_shards & (n_shards - 1)) || rank < 0 ||
        rank >= n_shards || n_players < 1) {
        ts_set_error("ts_rate_stream_sharded: bad argument (n_shards must be a power of two <= %d)", kMaxShards);
        return TS_E_ARG;
    }
    TS_CUDA(cudaSetDevice(ctx->device));
    cudaStream_t s = (cudaStream_t)stream;
    const int64_t rows = (n_players - rank + n_shards - 1) / n_shards;     // players p with p % G == rank
    unsigned long long *own = (unsigned long long *)vt_shards[rank];
    if (phase == 0 || phase == 2) {
        if (!table_shard || !own) { ts_set_error("ts_rate_stream_sharded: null shard"); return TS_E_ARG; }
        if (rows > 0) {
            if (phase == 0) k_shard_load<<<(unsigned)((rows + 255) / 256), 256, 0, s>>>(table_shard, own, rows);
            else k_shard_store<<<(unsigned)((rows + 255) / 256), 256, 0, s>>>(own, table_shard, rows);
        }
        TS_CUDA(cudaGetLastError());
        return 0;
    }
    if (phase != 1) { ts_set_error("ts_rate_stream_sharded: phase must be 0, 1 or 2"); return TS_E_ARG; }
    if (slots != 6 && slots != 10) { ts_set_error("ts_rate_stream_sharded: slots must be 6 or 10"); return TS_E_UNSUPPORTED; }
    if (n_matches < 0 || n_matches >= 0xFFFFFFFFll || (n_matches > 0 && (!ids || !meta || !pred))) {
        ts_set_error("ts_rate_stream_sharded: bad match arrays");
        return TS_E_ARG;
    }
    if (n_matches == 0) return 0;
    StreamArgs a;
    a.prm = ctx->params;
    a.vt = nullptr;
    a.seed = nullptr;
    for (int i = 0; i < 16; ++i) a.modes.col[i] = (int8_t)-1;
    if (mode >= 0 && mode < 6) a.modes.col[mode] = 0;
    a.ids = ids;
    a.meta = meta;
    a.pred = pred;
    a.n_matches = (uint32_t)n_matches;
    a.seq_base = 0;
    a.sleep_ns = stream_sleep_ns();
    a.max_spins = stream_max_spins();
    if (outputs) a.out = *outputs;
    else a.out = ts_outputs{nullptr, nullptr, nullptr, nullptr, nullptr};
    a.ctrl = ctx->d_ctrl;
    a.n_shards = (uint32_t)n_shards;
    a.shard_shift = 0;
    while ((1 << a.shard_shift) < n_shards) ++a.shard_shift;
    a.rank = (uint32_t)rank;
    for (int i = 0; i < kMaxShards; ++i) a.shards[i] = i < n_shards ? (unsigned long long *)vt_shards[i] : nullptr;
    for (int i = 0; i < n_shards; ++i)
        if (!a.shards[i]) { ts_set_error("ts_rate_stream_sharded: shard %d not mapped", i); return TS_E_ARG; }
    TS_CUDA(cudaMemsetAsync(ctx->d_ctrl + CTRL_STALLED, 0, sizeof(unsigned long long), s));
    int lanes = lanes_per_match ? lanes_per_match : (slots == 6 ? 2 : 8);   // measured: DESIGN.md 3.2
    return slots == 6 ? dispatch_l<6, true>(ctx, a, lanes, s) : dispatch_l<10, true>(ctx, a, lanes, s);
}
