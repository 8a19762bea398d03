// One ladder over the G GPUs of a node: the rating kernel of the sharded form
// (tsrate.h: ts_shard_view; the links come from ts_shard_link in sched.cu).
//
// Same dataflow rating as stream.cu -- matches in stream order, a match waits for
// the rows of its players -- but rows are PUSHED to where they will be read:
// every GPU has a mailbox of versioned rows (one per player), a rated match stores
// each new row, tagged with the number of that player's NEXT match, into the
// mailbox of the GPU that rates that match, and a match polls only its own GPU's
// mailbox until every row carries the match's own number.  A hop between GPUs is
// therefore one one-way NVLink store plus a local L2 poll; nothing ever polls
// across NVLink.  After a player's last match the pair goes to the owner's table
// shard (player p: GPU p % G, row p / G).  The tags are the only synchronisation
// between GPUs: no barrier, no collective, no fence (every 8-byte word of a row
// carries the tag, and 64-bit accesses are single-copy atomic).
//
// Work split: blocks of 32 consecutive matches go round-robin to the GPUs, so
// every GPU holds 1/G of the records and neighbouring blocks run in parallel.
// Progress: every GPU's grid is co-resident and takes its tiles in increasing
// order, so by induction on the match number the oldest unfinished match of the
// ladder never waits on work that has not started.
//
// EMU: one launch plays all G ranks (block b is a block of rank b % G; every
// pointer of the view is local).  Same code path, one GPU: this is what the
// single-GPU tests and the driver's one-GPU box run.
#include "stream_common.cuh"

namespace {

using namespace tsstream;

constexpr int kBlock = 256;

struct PushArgs {
    ts_params prm;
    ts_shard_view view;
    uint32_t g_mask, g_shift;
    uint32_t rated_mask;
    uint32_t n_matches;           // of the window, all ranks together
    uint32_t n_players;
    uint32_t seq_base;
    uint32_t sleep_ns, max_spins;
    uint32_t prefetch;
    ts_outputs out;
    unsigned long long *ctrl;
};

__device__ __forceinline__ void st_pair_sys(double *p, double mu, double sigma)
{
    asm volatile("st.relaxed.sys.global.v2.f64 [%0], {%1, %2};" ::"l"(p), "d"(mu), "d"(sigma) : "memory");
}

// matches of the window that live on rank r
__device__ __forceinline__ uint32_t local_matches(uint32_t n, uint32_t r, uint32_t g_mask, uint32_t g_shift)
{
    const uint32_t blocks = (n + 31u) >> 5;
    if (blocks <= r) return 0u;
    const uint32_t mine = (blocks - r + g_mask) >> g_shift;
    const bool last_is_mine = ((blocks - 1u) & g_mask) == r;
    return (mine - 1u) * 32u + (last_is_mine ? n - (blocks - 1u) * 32u : 32u);
}

template <int P, int L, bool EMU, bool HOIST>
__global__ void __launch_bounds__(kBlock, 2) k_rate_push(const PushArgs a)
{
    constexpr int W = 32 / L;
    constexpr int SPL = (P + L - 1) / L;
    constexpr int kWarps = kBlock / 32;
    __shared__ __align__(16) WarpStage<P, W> stage_all[kWarps];

    const int wl = threadIdx.x & 31;
    const int lane = wl % L;
    const int grp = wl / L;
    WarpStage<P, W> &stage = stage_all[threadIdx.x >> 5];
    const uint32_t me = EMU ? (blockIdx.x & a.g_mask) : (uint32_t)a.view.rank;
    const uint32_t block_of_rank = EMU ? (blockIdx.x >> a.g_shift) : blockIdx.x;
    const uint32_t blocks_of_rank = EMU ? (gridDim.x >> a.g_shift) : gridDim.x;
    const uint32_t *ids = a.view.ids[me], *meta_arr = a.view.meta[me], *succ = a.view.succ[me];
    unsigned long long *rows = a.view.rows[me];
    PassEnv env;
    env.tau2 = a.prm.tau * a.prm.tau;
    env.beta2 = a.prm.beta * a.prm.beta;
    env.tb2 = env.tau2 + env.beta2;
    env.margin_full = a.prm.draw_margin[P];
    env.margins = a.prm.draw_margin;
    env.seed = nullptr;
    env.draw_is_zero = a.prm.draw_is_zero != 0;
    const uint32_t n_local = local_matches(a.n_matches, me, a.g_mask, a.g_shift);
    const uint32_t n_tiles = (n_local + W - 1) / W;
    const uint32_t stride = blocks_of_rank * kWarps;
    const uint32_t first_tile = (threadIdx.x >> 5) * blocks_of_rank + block_of_rank;

    const unsigned long long pol = stream_policy();
    auto stage_tile = [&](int b, uint32_t t) {
        if (t < n_tiles) stage_warp_tile<P, W>(stage, b, ids, succ, meta_arr, t * W, min((uint32_t)W, n_local - t * W), wl, pol);
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
        if (D == 3 && a.prefetch && tile + stride < n_tiles)    // next tile's mailbox rows on their way into L2
            prefetch_tile_rows<P, W>(stage, nxt, rows, 4u, min((uint32_t)W, n_local - (tile + stride) * W), a.n_players, wl);
        const uint32_t base = tile * W;
        const uint32_t cnt = min((uint32_t)W, n_local - base);
        const uint32_t lm = base + grp;                                  // local index on this rank
        const uint32_t m = ((((lm >> 5) << a.g_shift) | me) << 5) | (lm & 31u);   // number in the window
        const uint32_t tag = a.seq_base + m;                             // stream-wide number: the tag to wait for
        const uint32_t om = EMU ? m : lm;                                // where this match's outputs go
        const uint32_t meta = (uint32_t)grp < cnt ? stage.meta[cur][grp] : (15u << TS_META_MODE_SHIFT);
        const uint32_t mode = (meta >> TS_META_MODE_SHIFT) & 15u;
        const bool skip_mode = ((a.rated_mask >> mode) & 1u) == 0u;
        const bool gated = skip_mode || (meta & (TS_META_AFK | TS_META_INVALID)) != 0u;
        const bool live = (uint32_t)grp < cnt && !gated;
        if ((uint32_t)grp < cnt && gated && lane == 0) {     // rater.py:83-85,102-106: outputs only
            if (a.out.status) a.out.status[om] = skip_mode ? TS_ST_SKIP_MODE : ((meta & TS_META_INVALID) ? TS_ST_INVALID : TS_ST_AFK);
            if (!skip_mode && a.out.quality) a.out.quality[om] = 0.0;
        }
        const int rank0 = (meta & TS_META_WIN_A) ? 0 : 1, rank1 = (meta & TS_META_WIN_B) ? 0 : 1;
        const int first = rank1 < rank0 ? 1 : 0;
        const bool is_draw = rank0 == rank1;

        uint32_t pid[SPL], next[SPL], want[SPL];
        bool used[SPL], have[SPL];
        double q_mu[SPL], q_sg[SPL];
        bool bad_id = false;
#pragma unroll
        for (int k = 0; k < SPL; ++k) {
            const int s = lane + k * L;
            pid[k] = (live && s < P) ? stage.ids[cur][grp * P + s] : TS_EMPTY_SLOT;
            used[k] = pid[k] != TS_EMPTY_SLOT;
            if (used[k] && pid[k] >= a.n_players) { used[k] = false; bad_id = true; }   // corrupt id: nobody links it
            next[k] = used[k] ? stage.link[cur][grp * P + s] : TS_NONE;
            want[k] = tag;
            have[k] = !used[k];
            q_mu[k] = 0.0; q_sg[k] = 1.0;
        }
        bool done = !live;
        uint32_t spins = 0;
        // wait loop and converged-warp pass as in stream.cu; every poll is a load from this
        // GPU's own mailbox (system scope: the writer may be another GPU)
        while (!__all_sync(0xFFFFFFFFu, done)) {
            bool ready = false;
            if (!done) ready = poll_rows<SPL, HOIST, true>(rows, pid, want, have, q_mu, q_sg);
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
                    if (next[k] != TS_NONE) {
                        // to the mailbox of the GPU that rates the player's next match, under that match's number
                        const uint32_t to = ((next[k] - a.seq_base) >> 5) & a.g_mask;
                        st_row<true>(a.view.rows[to] + (size_t)pid[k] * 4, mu2[k], sg2[k], next[k]);
                    } else {
                        // the player's last match of the window: back to the owner's table shard
                        st_pair_sys(a.view.table[pid[k] & a.g_mask] + (size_t)(pid[k] >> a.g_shift) * 2, mu2[k], sg2[k]);
                    }
                    if (ok && a.out.out_queue) st_pair_stream(a.out.out_queue + ((size_t)om * P + s) * 2, mu2[k], sg2[k], pol);
                }
            }
            if (go) {
                if (lane == 0) {
                    if (a.out.status) a.out.status[om] = (uint8_t)status;
                    if (!ok) atomicAdd(&a.ctrl[status == TS_ST_VALUE_ERROR ? CTRL_VALUE_ERRORS : CTRL_ERRORS], 1ull);
                    else if (a.out.quality) a.out.quality[om] = match_quality(d_q, ss_q, n, env.beta2);
                }
                done = true;
            }
        }
    }
}

template <int P, int L>
int launch_push(ts_ctx *ctx, const PushArgs &args, bool emulate, bool hoist, cudaStream_t stream)
{
    void *fn = emulate ? (hoist ? (void *)k_rate_push<P, L, true, true> : (void *)k_rate_push<P, L, true, false>)
                       : (hoist ? (void *)k_rate_push<P, L, false, true> : (void *)k_rate_push<P, L, false, false>);
    int per_sm = 0;
    TS_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, fn, kBlock, 0));
    if (per_sm < 1) { ts_set_error("sharded stream kernel does not fit on an SM"); return TS_E_CUDA; }
    per_sm = per_sm < stream_blocks_per_sm_cap() ? per_sm : stream_blocks_per_sm_cap();
    unsigned blocks = (unsigned)(ctx->sm_count * per_sm);
    if (emulate) blocks &= ~(unsigned)args.g_mask;         // the same number of blocks for every played rank
    if (blocks == 0) { ts_set_error("ts_shard_rate: more ranks to play than resident blocks"); return TS_E_UNSUPPORTED; }
    dim3 grid(blocks), block(kBlock);
    void *kargs[] = {(void *)&args};
    // cooperative launch only for its guarantee that the whole grid is resident
    TS_CUDA(cudaLaunchCooperativeKernel(fn, grid, block, kargs, 0, stream));
    return 0;
}

template <int P>
int dispatch_push(ts_ctx *ctx, const PushArgs &a, int lanes, bool emulate, bool hoist, cudaStream_t s)
{
    switch (lanes) {
    case 2: return launch_push<P, 2>(ctx, a, emulate, hoist, s);
    case 4: return launch_push<P, 4>(ctx, a, emulate, hoist, s);
    case 8: return launch_push<P, 8>(ctx, a, emulate, hoist, s);
    }
    ts_set_error("ts_shard_rate: lanes_per_match must be 0, 2, 4 or 8 (got %d)", lanes);
    return TS_E_ARG;
}

int check_shards(const ts_shard_view *v, const char *who)
{
    if (!v || v->n_shards < 1 || v->n_shards > TS_SHARD_MAX || (v->n_shards & (v->n_shards - 1)) || v->rank < 0 ||
        v->rank >= v->n_shards) {
        ts_set_error("%s: n_shards must be a power of two <= %d and 0 <= rank < n_shards", who, TS_SHARD_MAX);
        return TS_E_ARG;
    }
    return 0;
}

}  // namespace

int ts_shard_apply_links(const ts_shard_view *view, int me, int64_t n_matches, int32_t slots, cudaStream_t s);   // sched.cu

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

// Every mailbox row := tag TS_NONE (no match carries that number), and the sticky flags of
// this context are cleared.  The ranks meet at a barrier before anyone links.
extern "C" int ts_shard_reset(ts_ctx *ctx, const ts_shard_view *view, int64_t n_players, void *stream)
{
    if (int rc = check_shards(view, "ts_shard_reset")) return rc;
    if (!ctx || n_players < 1 || !view->rows[view->rank]) { ts_set_error("ts_shard_reset: bad argument"); return TS_E_ARG; }
    TS_CUDA(cudaSetDevice(ctx->device));
    cudaStream_t s = (cudaStream_t)stream;
    TS_CUDA(cudaMemsetAsync(view->rows[view->rank], 0xFF, (size_t)n_players * 32, s));
    TS_CUDA(cudaMemsetAsync(ctx->d_ctrl + CTRL_STALLED, 0, sizeof(unsigned long long), s));
    TS_CUDA(cudaMemsetAsync(ctx->d_ctrl + CTRL_SHARD, 0, sizeof(unsigned long long), s));
    return 0;
}

extern "C" int ts_shard_status(ts_ctx *ctx, void *stream)
{
    if (!ctx) return TS_E_ARG;
    unsigned long long stalled = 0, bits = 0;
    cudaStream_t s = (cudaStream_t)stream;
    TS_CUDA(cudaMemcpyAsync(&stalled, ctx->d_ctrl + CTRL_STALLED, sizeof stalled, cudaMemcpyDeviceToHost, s));
    TS_CUDA(cudaMemcpyAsync(&bits, ctx->d_ctrl + CTRL_SHARD, sizeof bits, cudaMemcpyDeviceToHost, s));
    TS_CUDA(cudaStreamSynchronize(s));
    return (int)((bits & 6ull) | (stalled ? (unsigned long long)TS_SHARD_STALLED : 0ull));
}

extern "C" int ts_shard_rate(ts_ctx *ctx, const ts_shard_view *view, int64_t n_matches, int32_t slots, int64_t n_players,
                             int8_t mode, int64_t seq_base, const ts_outputs *outputs, int32_t lanes_per_match,
                             int32_t emulate_all, void *stream)
{
    if (int rc = check_shards(view, "ts_shard_rate")) return rc;
    const int G = view->n_shards;
    if (!ctx || n_players < 1 || n_players >= 0xFFFFFFFFll || mode < 0 || mode > 5) {
        ts_set_error("ts_shard_rate: bad argument");
        return TS_E_ARG;
    }
    if (slots != 6 && slots != 10) { ts_set_error("ts_shard_rate: slots must be 6 or 10"); return TS_E_UNSUPPORTED; }
    if (n_matches < 0 || seq_base < 0 || seq_base % (TS_SHARD_BLOCK * G) != 0 || seq_base + n_matches >= 0xFFFFFFFFll) {
        ts_set_error("ts_shard_rate: seq_base must be a multiple of 32 * n_shards and match numbers stay below 2^32 - 1");
        return TS_E_ARG;
    }
    if (n_matches == 0) return 0;
    for (int g = 0; g < G; ++g) {
        const bool need_records = emulate_all || g == view->rank;
        if (!view->rows[g] || !view->table[g] || (need_records && (!view->ids[g] || !view->meta[g] || !view->succ[g]))) {
            ts_set_error("ts_shard_rate: shard %d not mapped", g);
            return TS_E_ARG;
        }
    }
    TS_CUDA(cudaSetDevice(ctx->device));
    PushArgs a;
    a.prm = ctx->params;
    a.view = *view;
    a.g_mask = (uint32_t)G - 1u;
    a.g_shift = 0;
    while ((1 << a.g_shift) < G) ++a.g_shift;
    a.rated_mask = 1u << mode;
    a.n_matches = (uint32_t)n_matches;
    a.n_players = (uint32_t)n_players;
    a.seq_base = (uint32_t)seq_base;
    a.sleep_ns = stream_sleep_ns();
    a.max_spins = stream_max_spins();
    if (outputs) a.out = *outputs;
    else a.out = ts_outputs{nullptr, nullptr, nullptr, nullptr, nullptr};
    a.ctrl = ctx->d_ctrl;
    const int lanes = lanes_per_match ? lanes_per_match : (slots == 6 ? 2 : 8);   // measured: DESIGN.md 3.2
    cudaStream_t s = (cudaStream_t)stream;
    // the links every sender left in the inbox(es) go to their slots first: local work, same stream
    for (int g = 0; g < G; ++g) {
        if (!(emulate_all || g == view->rank)) continue;
        if (!view->inbox[g] || !view->inbox_count[g] || view->inbox_cap < 1) {
            ts_set_error("ts_shard_rate: inbox of shard %d not mapped", g);
            return TS_E_ARG;
        }
        if (int rc = ts_shard_apply_links(view, g, n_matches, slots, s)) return rc;
    }
    const bool hoist = stream_hoist_mode() != 0;
    a.prefetch = stream_prefetch_mode() != 0 ? 1u : 0u;
    return slots == 6 ? dispatch_push<6>(ctx, a, lanes, emulate_all != 0, hoist, s)
                      : dispatch_push<10>(ctx, a, lanes, emulate_all != 0, hoist, s);
}
