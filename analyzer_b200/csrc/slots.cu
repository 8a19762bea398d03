// Streaming (dataflow) rating with rows indexed by the CONSUMING SLOT -- for ladders whose
// rows do not fit L2.
//
// ts_rate_stream keeps one versioned row per player; where the 32 bytes x players exceed L2
// every poll is a scattered DRAM sector read and every publish a scattered sector write (ncu,
// 125M 3v3 matches over 12.5M players: 104 GB of DRAM traffic, 3.8x the algorithmic bytes, 55 %
// of the copy bandwidth).  Here a row lives where it will be READ: rows[m * P + s] is the rating
// the player in slot s of match m brings to that match.  The match that produced it stored it
// there (succ[v] names the slot of the player's next match; ts_link_succ computes it from the
// same sorted runs as pred), so a match reads its P rows as P * 32 consecutive bytes -- the
// reads of a whole warp tile are one contiguous piece -- and only the writes scatter (full
// sectors, no read-modify-write).  A player's first row of the window comes from the table
// (written by the link kernel), the last one goes back to the table: no separate load / store
// pass over the table, and players that do not play are never touched.
//
// A row is valid when its four words carry the EPOCH of this call (every word validates itself,
// as in stream.cu); the caller passes a different epoch each time it reuses the rows buffer, so
// the buffer is never cleared (zero it once when it is allocated).  Same pass (rate_pass), same
// bits as ts_rate_stream.  Cost: 32 bytes of workspace per slot of the window instead of per
// player; no resident state between windows (the table carries it).
#include "stream_common.cuh"

namespace {

using namespace tsstream;

constexpr int kBlock = 256;

struct SlotArgs {
    ts_params prm;
    double *table;                // [U][2]
    const double *seed;
    unsigned long long *rows;     // [n_matches * P] rows of four words
    const uint32_t *ids;
    const uint32_t *meta;
    const uint32_t *succ;
    uint32_t rated_mask;
    uint32_t n_matches;
    uint32_t n_players;
    uint32_t epoch;
    uint32_t sleep_ns, max_spins;
    ts_outputs out;
    unsigned long long *ctrl;
};

template <int P, int L>
__global__ void __launch_bounds__(kBlock, 2) k_rate_slots(const SlotArgs a)
{
    constexpr int W = 32 / L;
    constexpr int SPL = (P + L - 1) / L;
    constexpr int kWarps = kBlock / 32;
    constexpr int D = WarpStage<P, W>::kDepth;
    __shared__ __align__(16) WarpStage<P, W> stage_all[kWarps];

    const int wl = threadIdx.x & 31;
    const int lane = wl % L;
    const int grp = wl / L;
    WarpStage<P, W> &stage = stage_all[threadIdx.x >> 5];
    PassEnv env;
    env.tau2 = a.prm.tau * a.prm.tau;
    env.beta2 = a.prm.beta * a.prm.beta;
    env.tb2 = env.tau2 + env.beta2;
    env.margin_full = a.prm.draw_margin[P];
    env.margins = a.prm.draw_margin;
    env.seed = a.seed;
    env.draw_is_zero = a.prm.draw_is_zero;
    const unsigned long long pol = stream_policy();
    const uint32_t n_tiles = (a.n_matches + W - 1) / W;
    const uint32_t stride = gridDim.x * kWarps;
    const uint32_t first_tile = (threadIdx.x >> 5) * gridDim.x + blockIdx.x;
    auto stage_tile = [&](int b, uint32_t t) {
        if (t < n_tiles) stage_warp_tile<P, W>(stage, b, a.ids, a.succ, a.meta, t * W, min((uint32_t)W, a.n_matches - t * W), wl, pol);
    };
    int cur = 0;
    stage_tile(0, first_tile);
    if (D == 3) stage_tile(1, first_tile + stride);
    for (uint32_t tile = first_tile; tile < n_tiles; tile += stride) {
        cp_async_wait_all();
        __syncwarp();
        const int nxt = cur == D - 1 ? 0 : cur + 1;
        stage_tile((cur + D - 1) % D, tile + (D - 1) * stride);
        const uint32_t base = tile * W;
        const uint32_t cnt = min((uint32_t)W, a.n_matches - base);
        const uint32_t m = base + grp;
        const uint32_t meta = (uint32_t)grp < cnt ? stage.meta[cur][grp] : (15u << TS_META_MODE_SHIFT);
        const uint32_t mode = (meta >> TS_META_MODE_SHIFT) & 15u;
        const bool skip_mode = ((a.rated_mask >> mode) & 1u) == 0u;
        const bool gated = skip_mode || (meta & (TS_META_AFK | TS_META_INVALID)) != 0u;
        const bool live = (uint32_t)grp < cnt && !gated;
        if ((uint32_t)grp < cnt && gated && lane == 0) {     // rater.py:83-85,102-106: outputs only
            if (a.out.status) a.out.status[m] = skip_mode ? TS_ST_SKIP_MODE : ((meta & TS_META_INVALID) ? TS_ST_INVALID : TS_ST_AFK);
            if (!skip_mode && a.out.quality) a.out.quality[m] = 0.0;
        }
        const int rank0 = (meta & TS_META_WIN_A) ? 0 : 1, rank1 = (meta & TS_META_WIN_B) ? 0 : 1;
        const int first = rank1 < rank0 ? 1 : 0;
        const bool is_draw = rank0 == rank1;

        uint32_t pid[SPL], slot[SPL], next[SPL], want[SPL];
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
            slot[k] = m * P + (s < P ? s : 0);
            want[k] = a.epoch;
            have[k] = !used[k];
            q_mu[k] = 0.0; q_sg[k] = 1.0;
        }
        cur = nxt;
        bool done = !live;
        uint32_t spins = 0;
        while (!__all_sync(0xFFFFFFFFu, done)) {
            bool ready = false;
            if (!done) ready = poll_rows<SPL, true, false>(a.rows, slot, want, have, q_mu, q_sg);
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
                    // to the slot of the player's next match, or -- last match of the window -- to the table
                    if (next[k] != TS_NONE) st_row(a.rows + (size_t)next[k] * 4, mu2[k], sg2[k], a.epoch);
                    else __stcg(reinterpret_cast<double2 *>(a.table) + pid[k], make_double2(mu2[k], sg2[k]));
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

template <int P, int L>
int launch_slots(ts_ctx *ctx, const SlotArgs &args, cudaStream_t stream)
{
    int per_sm = 0;
    TS_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_rate_slots<P, L>, kBlock, 0));
    if (per_sm < 1) { ts_set_error("slot stream kernel does not fit on an SM"); return TS_E_CUDA; }
    per_sm = per_sm < stream_blocks_per_sm_cap() ? per_sm : stream_blocks_per_sm_cap();
    dim3 grid(ctx->sm_count * per_sm), block(kBlock);
    void *kargs[] = {(void *)&args};
    // cooperative launch only for its guarantee that the whole grid is resident
    TS_CUDA(cudaLaunchCooperativeKernel((void *)k_rate_slots<P, L>, grid, block, kargs, 0, stream));
    return 0;
}

}  // namespace

extern "C" int64_t ts_slot_rows_bytes(int64_t n_matches, int32_t slots)
{
    if (n_matches < 0 || slots < 2 || slots > TS_MAX_SLOTS) return TS_E_ARG;
    return n_matches * slots * 32;
}

extern "C" int ts_rate_slots(ts_ctx *ctx, const ts_table *table, const uint32_t *ids, const uint32_t *meta,
                             const uint32_t *succ, int64_t n_matches, int32_t slots, uint32_t epoch,
                             const ts_outputs *outputs, int32_t lanes_per_match, void *rows, void *stream)
{
    if (!ctx || !table || !table->data || (n_matches > 0 && (!ids || !meta || !succ || !rows))) {
        ts_set_error("ts_rate_slots: null argument");
        return TS_E_ARG;
    }
    if (table->row_pairs != 1 || table->shared_col >= 0) {
        ts_set_error("ts_rate_slots: the table must be a single-pair ladder (row_pairs == 1, shared_col == -1)");
        return TS_E_UNSUPPORTED;
    }
    if (slots != 6 && slots != 10) { ts_set_error("ts_rate_slots: slots must be 6 or 10 (got %d)", slots); return TS_E_UNSUPPORTED; }
    if (n_matches < 0 || (uint64_t)n_matches * (uint64_t)slots >= 0xFFFF0000ull) {
        ts_set_error("ts_rate_slots: n_matches * slots must stay below 2^32 - 2^16: rate the ladder in windows");
        return TS_E_ARG;
    }
    if (epoch == 0u || epoch == TS_NONE || ((uintptr_t)rows & 31u) != 0u) {
        ts_set_error("ts_rate_slots: epoch must be neither 0 nor 0xFFFFFFFF, rows 32-byte aligned");
        return TS_E_ARG;
    }
    if (n_matches == 0) return 0;
    TS_CUDA(cudaSetDevice(ctx->device));
    SlotArgs a;
    a.prm = ctx->params;
    a.table = table->data;
    a.seed = table->seed;
    a.rows = (unsigned long long *)rows;
    a.ids = ids;
    a.meta = meta;
    a.succ = succ;
    a.rated_mask = 0;
    for (int i = 0; i < 6; ++i) if (table->mode_col[i] >= 0) a.rated_mask |= 1u << i;
    a.n_matches = (uint32_t)n_matches;
    a.n_players = (uint32_t)(table->n_players > 0xFFFFFFFFll ? 0xFFFFFFFFll : table->n_players);
    a.epoch = epoch;
    a.sleep_ns = stream_sleep_ns();
    a.max_spins = stream_max_spins();
    if (outputs) a.out = *outputs;
    else a.out = ts_outputs{nullptr, nullptr, nullptr, nullptr, nullptr};
    a.ctrl = ctx->d_ctrl;
    cudaStream_t s = (cudaStream_t)stream;
    const int lanes = lanes_per_match ? lanes_per_match : (slots == 6 ? 2 : 8);   // measured: DESIGN.md 3.2
    if (slots == 6) {
        switch (lanes) {
        case 2: return launch_slots<6, 2>(ctx, a, s);
        case 4: return launch_slots<6, 4>(ctx, a, s);
        case 8: return launch_slots<6, 8>(ctx, a, s);
        }
    } else {
        switch (lanes) {
        case 2: return launch_slots<10, 2>(ctx, a, s);
        case 4: return launch_slots<10, 4>(ctx, a, s);
        case 8: return launch_slots<10, 8>(ctx, a, s);
        }
    }
    ts_set_error("ts_rate_slots: lanes_per_match must be 0, 2, 4 or 8 (got %d)", lanes);
    return TS_E_ARG;
}
