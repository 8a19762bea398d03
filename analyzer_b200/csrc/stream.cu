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

struct StreamArgs {
    ts_params prm;
    unsigned long long *vt;       // versioned rows, 4 words per player
    const double *seed;
    ModeCols modes;
    const uint32_t *ids;
    const uint32_t *meta;
    const uint32_t *pred;
    uint32_t n_matches;
    uint32_t n_players;           // ids at or beyond it are refused (TS_ST_VALUE_ERROR), never dereferenced
    uint32_t seq_base;            // stream-wide number of the first match of this window
    uint32_t sleep_ns;            // back-off between polls when no group of the warp is complete
    uint32_t max_spins;           // watchdog: fruitless polls before a warp declares the stream stalled
    uint32_t prefetch;            // 1: the rows of the next tile's players are prefetched into L2
    ts_outputs out;
    unsigned long long *ctrl;
};

// Warps are the unit of work: warp w of the grid takes the tiles w, w + n_warps,
// ... of W consecutive matches, in increasing order, and never synchronises
// with the other warps of its block -- a warp that has to wait for a dependency
// holds up nothing but its own W matches.
template <int P, int L, bool HOIST>
__global__ void __launch_bounds__(kBlock, TS_STREAM_MIN_BLOCKS) k_rate_stream(const StreamArgs a)
{
    constexpr int W = 32 / L;
    constexpr int SPL = (P + L - 1) / L;
    constexpr int kWarps = kBlock / 32;
    __shared__ __align__(16) WarpStage<P, W> stage_all[kWarps];

    const int wl = threadIdx.x & 31;            // lane in warp
    const int lane = wl % L;                    // lane in match group
    const int grp = wl / L;                     // match of the warp tile
    WarpStage<P, W> &stage = stage_all[threadIdx.x >> 5];
    PassEnv env;
    env.tau2 = a.prm.tau * a.prm.tau;
    env.beta2 = a.prm.beta * a.prm.beta;
    env.tb2 = env.tau2 + env.beta2;
    env.margin_full = a.prm.draw_margin[P];
    env.margins = a.prm.draw_margin;
    env.seed = a.seed;
    env.draw_is_zero = a.prm.draw_is_zero;
    uint32_t rated_mask = 0;
#pragma unroll
    for (int i = 0; i < 6; ++i) rated_mask |= a.modes.col[i] >= 0 ? (1u << i) : 0u;
    const unsigned long long pol = stream_policy();
    const uint32_t n_tiles = (a.n_matches + W - 1) / W;
    const uint32_t stride = gridDim.x * kWarps;
    // neighbouring tiles go to different SMs
    const uint32_t first_tile = (threadIdx.x >> 5) * gridDim.x + blockIdx.x;
    auto stage_tile = [&](int b, uint32_t t) {
        if (t < n_tiles) stage_warp_tile<P, W>(stage, b, a.ids, a.pred, a.meta, t * W, min((uint32_t)W, a.n_matches - t * W), wl, pol);
    };

    int cur = 0;
    stage_tile(0, first_tile);
    constexpr int D = WarpStage<P, W>::kDepth;
    if (D == 3) stage_tile(1, first_tile + stride);
    for (uint32_t tile = first_tile; tile < n_tiles; tile += stride) {
        cp_async_wait_all();                    // this tile's records and the next one's (issued a whole tile ago)
        __syncwarp();
        const int nxt = cur == D - 1 ? 0 : cur + 1;
        stage_tile((cur + D - 1) % D, tile + (D - 1) * stride);
        if (D == 3 && a.prefetch && tile + stride < n_tiles)    // next tile's rows on their way into L2 while this one waits and computes
            prefetch_tile_rows<P, W>(stage, nxt, a.vt, 4u, min((uint32_t)W, a.n_matches - (tile + stride) * W), a.n_players, wl);
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
        bool bad_id = false;
#pragma unroll
        for (int k = 0; k < SPL; ++k) {
            const int s = lane + k * L;
            pid[k] = (live && s < P) ? stage.ids[cur][grp * P + s] : TS_EMPTY_SLOT;
            used[k] = pid[k] != TS_EMPTY_SLOT;
            if (used[k] && pid[k] >= a.n_players) { used[k] = false; bad_id = true; }   // corrupt id: no table access
            want[k] = used[k] ? stage.link[cur][grp * P + s] : TS_NONE;
            have[k] = !used[k];
            q_mu[k] = 0.0; q_sg[k] = 1.0;
        }
        cur = nxt;
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
            if (!done) ready = poll_rows<SPL, HOIST, false>(a.vt, pid, want, have, q_mu, q_sg);
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

            double mu2[SPL], sg2[SPL], d_q, ss_q;
            int n;
            const int status = rate_pass<P, L, SPL>(env, lane, first, draw_eff, pid, used, bad_id, q_mu, q_sg, mu2, sg2,
                                                    d_q, ss_q, n);
            const bool ok = status == TS_ST_OK;
#pragma unroll
            for (int k = 0; k < SPL; ++k) {
                const int s = lane + k * L;
                if (go && used[k]) {
                    st_row(a.vt + (size_t)pid[k] * 4, mu2[k], sg2[k], a.seq_base + m);   // first: successors are waiting
                    if (ok && a.out.out_queue) st_pair_stream(a.out.out_queue + ((size_t)m * P + s) * 2, mu2[k], sg2[k], pol);
                }
            }
            if (go) {
                if (lane == 0) {
                    if (a.out.status) a.out.status[m] = (uint8_t)status;
                    if (!ok) atomicAdd(&a.ctrl[status == TS_ST_VALUE_ERROR ? CTRL_VALUE_ERRORS : CTRL_ERRORS], 1ull);
                    else if (a.out.quality) a.out.quality[m] = match_quality(d_q, ss_q, n, env.beta2);
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

template <int P, int L>
int launch_stream(ts_ctx *ctx, const StreamArgs &args, bool hoist, cudaStream_t stream)
{
    void *fn = hoist ? (void *)k_rate_stream<P, L, true> : (void *)k_rate_stream<P, L, false>;
    int per_sm = 0;
    TS_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, fn, kBlock, 0));
    if (per_sm < 1) { ts_set_error("stream kernel does not fit on an SM"); return TS_E_CUDA; }
    per_sm = per_sm < stream_blocks_per_sm_cap() ? per_sm : stream_blocks_per_sm_cap();
    dim3 grid(ctx->sm_count * per_sm), block(kBlock);
    void *kargs[] = {(void *)&args};
    // cooperative launch only for its guarantee that the whole grid is resident
    TS_CUDA(cudaLaunchCooperativeKernel(fn, grid, block, kargs, 0, stream));
    return 0;
}

template <int P>
int dispatch_l(ts_ctx *ctx, const StreamArgs &a, int lanes, bool hoist, cudaStream_t s)
{
    switch (lanes) {
    case 1: return launch_stream<P, 1>(ctx, a, hoist, s);
    case 2: return launch_stream<P, 2>(ctx, a, hoist, s);
    case 4: return launch_stream<P, 4>(ctx, a, hoist, s);
    case 8: return launch_stream<P, 8>(ctx, a, hoist, s);
    case 16: return launch_stream<P, 16>(ctx, a, hoist, s);
    case 32: return launch_stream<P, 32>(ctx, a, hoist, s);
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
    if (((uintptr_t)workspace & 31u) != 0u) {
        ts_set_error("ts_rate_stream: workspace must be 32-byte aligned (rows are read and written as 256-bit words)");
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
    a.n_players = (uint32_t)(table->n_players > 0xFFFFFFFFll ? 0xFFFFFFFFll : table->n_players);
    a.seq_base = (uint32_t)seq_base;
    a.sleep_ns = stream_sleep_ns();
    a.max_spins = stream_max_spins();
    if (outputs) a.out = *outputs;
    else a.out = ts_outputs{nullptr, nullptr, nullptr, nullptr, nullptr};
    a.ctrl = ctx->d_ctrl;
    const int64_t u = table->n_players;
    if (flags & TS_STREAM_LOAD) k_vt_load<<<(unsigned)((u + 255) / 256), 256, 0, s>>>(table->data, a.vt, u);
    int lanes = lanes_per_match ? lanes_per_match : (slots == 6 ? 2 : 8);   // measured: DESIGN.md 3.2
    const bool hoist = stream_hoist_mode() != 0;
    a.prefetch = stream_prefetch_mode() != 0 ? 1u : 0u;
    int rc = 0;
    if (n_matches > 0) rc = slots == 6 ? dispatch_l<6>(ctx, a, lanes, hoist, s) : dispatch_l<10>(ctx, a, lanes, hoist, s);
    if (rc) return rc;
    if (flags & TS_STREAM_STORE) k_vt_store<<<(unsigned)((u + 255) / 256), 256, 0, s>>>(a.vt, table->data, u);
    TS_CUDA(cudaGetLastError());
    return 0;
}


