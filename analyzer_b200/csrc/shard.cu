// One ladder over the G GPUs of a node: the rating kernel of the sharded form
// (tsrate.h: ts_shard_view; the links come from ts_shard_link in sched.cu).
//
// Same dataflow rating as stream.cu -- matches in stream order, a match waits for
// the rows of its players -- but rows are PUSHED to where they will be read:
// every GPU has a mailbox of versioned rows (one per player), a rated match sends
// each new row, tagged with the number of that player's NEXT match, to the GPU that
// rates that match, and a match polls only its own GPU's mailbox until every row
// carries the match's own number.  Nothing ever polls across NVLink.  After a
// player's last match the pair goes back to the owner's table shard (player p:
// GPU p % G, row p / G).  The tags are the only synchronisation between GPUs: no
// barrier, no collective, no fence (every 8-byte word of a row carries the tag,
// and 64-bit accesses are single-copy atomic).
//
// How a row crosses NVLink: NOT as a store to its mailbox slot.  Stores scattered
// over several peers' memory collapse on B200 -- measured (scripts/nvlink_probe.py):
// 8.4 G random 32-byte stores/s into one peer's 320 MB, 0.14 G/s when they spread
// over seven peers, while stores to consecutive addresses keep 13 G/s (435 GB/s)
// towards seven peers.  So a row for another GPU is a 32-byte MESSAGE (the pair,
// the tag, the player) appended to this sender's ring in the receiver's memory --
// the lanes of a warp that send to the same GPU reserve consecutive places with one
// local atomic -- and a few ROUTER blocks of the receiver's own grid read their
// rings in order and put every message into its mailbox row (or table row, for a
// last match) with a local store.  A message validates itself like a row (a marker
// byte in each of its four words); the sender's final count, published by its last
// block, tells the routers when a ring is drained.
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
    uint32_t router_blocks;       // blocks of every rank's grid that route incoming messages (0 when there is one rank)
    uint32_t window_gen;          // number of this window + 1: what this launch's final counts carry
    uint32_t epoch;               // 1..255: the marker byte of this run's messages
    uint32_t direct;              // 1: rows go straight into the receiver's mailbox (no rings, no routers)
    uint32_t ring_cap;            // messages one sender may leave in one receiver's ring
    uint32_t stage_msgs;          // 1: a pass's messages are grouped by receiver in shared memory first (MsgStage)
    ts_outputs out;
    unsigned long long *ctrl;
};

__device__ __forceinline__ void st_pair_sys(double *p, double mu, double sigma)
{
    asm volatile("st.relaxed.sys.global.v2.f64 [%0], {%1, %2};" ::"l"(p), "d"(mu), "d"(sigma) : "memory");
}

// A message: four 64-bit words, each [marker:8 | 24 bits of (player, tag)][32 bits of (mu, sigma)].
// tag TS_NONE = the player's last match of the window: the pair goes to the owner's table shard.
// (the marker is the run's epoch, 1..255, counted by ts_shard_reset: a ring need not be cleared between runs)
// ring_tail section of a rank (own use only), 32-bit words: [0..15] places reserved so far in every
// receiver's ring (they grow over all windows of a run), [16] rating blocks done in this launch,
// [32..47] messages this rank has consumed from every sender's ring up to the end of the last window.
// A final count is a 64-bit word [window number + 1 | count], so that the one of an earlier window
// is never taken for this window's.
constexpr int kTailDone = TS_SHARD_MAX, kTailConsumed = 2 * TS_SHARD_MAX;

__device__ __forceinline__ void pack_message(uint32_t pid, uint32_t tag, double mu, double sigma, unsigned long long kMsgMarker,
                                             unsigned long long w[4])
{
    const unsigned long long m = (unsigned long long)__double_as_longlong(mu);
    const unsigned long long s = (unsigned long long)__double_as_longlong(sigma);
    const unsigned long long x0 = pid & 0xFFFFFFu, x1 = (pid >> 24) | ((tag & 0xFFFFu) << 8), x2 = tag >> 16;
    w[0] = (kMsgMarker << 56) | (x0 << 32) | (m & 0xFFFFFFFFull);
    w[1] = (kMsgMarker << 56) | (x1 << 32) | (m >> 32);
    w[2] = (kMsgMarker << 56) | (x2 << 32) | (s & 0xFFFFFFFFull);
    w[3] = (kMsgMarker << 56) | (s >> 32);
}

__device__ __forceinline__ void st_words_sys(unsigned long long *slot, const unsigned long long w[4])
{
    asm volatile("st.relaxed.sys.global.v4.u64 [%0], {%1, %2, %3, %4};" ::"l"(slot), "l"(w[0]), "l"(w[1]), "l"(w[2]), "l"(w[3]) : "memory");
}

__device__ __forceinline__ void st_message(unsigned long long *slot, uint32_t pid, uint32_t tag, double mu, double sigma,
                                           unsigned long long kMsgMarker)
{
    unsigned long long w[4];
    pack_message(pid, tag, mu, sigma, kMsgMarker, w);
    st_words_sys(slot, w);
}

// The messages of one pass of a warp, grouped by receiver before they leave: a warp's pass makes
// 32 * SPL rows for up to G receivers, and a run of messages for one receiver stored by
// neighbouring lanes crosses NVLink in fewer, larger packets than the same messages stored
// four at a time (the lanes of one store instruction that happen to share a receiver).
template <int N>
struct alignas(16) MsgStage {
    unsigned long long w[N][4];
    uint32_t at[N];
    uint8_t to[N];
};

// Router warp: drains one sender's ring in this rank's memory.  The warps that share a ring take its
// chunks of 32 * kRouteUnroll messages round-robin; a lane delivers each of its messages as soon as
// that message is there, whatever the rest of the chunk does.
constexpr int kRouteUnroll = 4;

__device__ void route_ring(const PushArgs &a, uint32_t me, uint32_t sender, uint32_t phase, uint32_t n_phases, int wl)
{
    const unsigned long long *ring = a.view.ring[me] + (size_t)sender * a.ring_cap * 4;
    const unsigned long long *final_word = a.view.ring_final[me] + sender;
    unsigned long long *rows = a.view.rows[me];
    double *table = a.view.table[me];
    uint32_t *consumed = a.view.ring_tail[me] + kTailConsumed + sender;
    const uint32_t start = *consumed;             // where the earlier windows of this run ended
    uint32_t total = 0xFFFFFFFFu;                 // unknown until the sender's last block has published it
    uint32_t spins = 0;
    for (uint32_t chunk = phase;; chunk += n_phases) {
        const uint32_t base = start + chunk * (32u * kRouteUnroll);
        bool got[kRouteUnroll];
#pragma unroll
        for (int u = 0; u < kRouteUnroll; ++u) got[u] = false;
        for (;;) {
            if (total == 0xFFFFFFFFu) {
                unsigned long long f;
                asm volatile("ld.relaxed.sys.global.u64 %0, [%1];" : "=l"(f) : "l"(final_word) : "memory");
                if ((uint32_t)(f >> 32) == a.window_gen) total = min((uint32_t)f, a.ring_cap);
            }
            if (base >= total) {                  // everything this warp is responsible for has been delivered
                if (phase == 0 && wl == 0) *consumed = total;      // the next window starts here (stream order: visible to the next launch)
                return;
            }
            unsigned long long w[kRouteUnroll][4];
#pragma unroll
            for (int u = 0; u < kRouteUnroll; ++u) {
                const uint32_t i = base + u * 32 + wl;
                if (!got[u] && i < total) ld_row<true>(ring + (size_t)i * 4, w[u]);
            }
            bool all = true;
#pragma unroll
            for (int u = 0; u < kRouteUnroll; ++u) {
                const uint32_t i = base + u * 32 + wl;
                if (got[u]) continue;
                if (i >= total) { if (total != 0xFFFFFFFFu) got[u] = true; else all = false; continue; }
                const unsigned long long kMsgMarker = a.epoch;
                if ((w[u][0] >> 56) == kMsgMarker && (w[u][1] >> 56) == kMsgMarker && (w[u][2] >> 56) == kMsgMarker &&
                    (w[u][3] >> 56) == kMsgMarker) {
                    const uint32_t x0 = (uint32_t)(w[u][0] >> 32) & 0xFFFFFFu, x1 = (uint32_t)(w[u][1] >> 32) & 0xFFFFFFu,
                                   x2 = (uint32_t)(w[u][2] >> 32) & 0xFFFFFFu;
                    const uint32_t pid = x0 | ((x1 & 0xFFu) << 24), tag = (x1 >> 8) | (x2 << 16);
                    const double mu = __longlong_as_double((long long)((w[u][0] & 0xFFFFFFFFull) | (w[u][1] << 32)));
                    const double sg = __longlong_as_double((long long)((w[u][2] & 0xFFFFFFFFull) | (w[u][3] << 32)));
                    if (tag != TS_NONE) st_row<true>(rows + (size_t)pid * 4, mu, sg, tag);
                    else st_pair_sys(table + (size_t)(pid >> a.g_shift) * 2, mu, sg);
                    got[u] = true;
                } else {
                    all = false;
                }
            }
            if (__all_sync(0xFFFFFFFFu, all)) break;
            if ((++spins & 0x3FFu) == 0u) {
                if (spins >= a.max_spins && wl == 0) atomicExch(&a.ctrl[CTRL_STALLED], 1ull);
                if (ld_relaxed_u64(&a.ctrl[CTRL_STALLED]) != 0ull) return;
            }
            __nanosleep(a.sleep_ns);
        }
    }
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
    // grouping of the outgoing messages where it fits beside the record staging (static shared memory, 48 KB)
    constexpr bool kCanStage = (sizeof(WarpStage<P, W>) + sizeof(MsgStage<32 * SPL>)) * kWarps <= 48 * 1024 - 64;
    __shared__ __align__(16) MsgStage<kCanStage ? 32 * SPL : 1> msgs_all[kCanStage ? kWarps : 1];

    const int wl = threadIdx.x & 31;
    const int lane = wl % L;
    const int grp = wl / L;
    WarpStage<P, W> &stage = stage_all[threadIdx.x >> 5];
    const uint32_t me = EMU ? (blockIdx.x & a.g_mask) : (uint32_t)a.view.rank;
    const uint32_t grid_block = EMU ? (blockIdx.x >> a.g_shift) : blockIdx.x;
    const uint32_t grid_blocks = EMU ? (gridDim.x >> a.g_shift) : gridDim.x;
    if (grid_block < a.router_blocks) {
        // router: warp rw of this rank's routers serves sender (rw % senders), phase rw / senders
        const uint32_t senders = a.g_mask, rw = grid_block * kWarps + (threadIdx.x >> 5);
        const uint32_t n_phases = (a.router_blocks * kWarps) / senders;
        if (rw >= n_phases * senders) return;
        const uint32_t si = rw % senders;
        route_ring(a, me, si < me ? si : si + 1u, rw / senders, n_phases, wl);
        return;
    }
    const uint32_t block_of_rank = grid_block - a.router_blocks;
    const uint32_t blocks_of_rank = grid_blocks - a.router_blocks;
    const uint32_t *ids = a.view.ids[me], *meta_arr = a.view.meta[me], *succ = a.view.succ[me];
    unsigned long long *rows = a.view.rows[me];
    PassEnv env;
    env.tau2 = a.prm.tau * a.prm.tau;
    env.beta2 = a.prm.beta * a.prm.beta;
    env.tb2 = env.tau2 + env.beta2;
    env.margin_full = a.prm.draw_margin[P];
    env.margins = a.prm.draw_margin;
    env.seed = nullptr;
    env.draw_is_zero = a.prm.draw_is_zero;
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
            // Where every new row goes: the GPU that rates the player's next match (under that match's
            // number), or the player's owner after the last match of the window.  Own GPU: a local store.
            // Another GPU: a message in its ring -- the lanes that send to the same GPU take consecutive
            // places, reserved by one of them with a local atomic (all reservations of a pass are issued
            // before any is used).
            uint32_t to[SPL], grp_mask[SPL], place[SPL];
#pragma unroll
            for (int k = 0; k < SPL; ++k) {
                const bool send = go && used[k];
                to[k] = !send ? 0xFFu : (next[k] != TS_NONE ? ((next[k] - a.seq_base) >> 5) & a.g_mask : pid[k] & a.g_mask);
                if (to[k] == me || (a.direct && to[k] != 0xFFu)) {
                    if (next[k] != TS_NONE) st_row<true>(a.view.rows[to[k]] + (size_t)pid[k] * 4, mu2[k], sg2[k], next[k]);
                    else st_pair_sys(a.view.table[to[k]] + (size_t)(pid[k] >> a.g_shift) * 2, mu2[k], sg2[k]);
                    to[k] = 0xFFu;
                }
            }
            if (kCanStage && a.stage_msgs) {
                if (!a.direct && a.g_mask != 0u) {
                    // lane d counts receiver d's messages of this pass and reserves their places with one atomic
                    auto &msgs = msgs_all[kCanStage ? (threadIdx.x >> 5) : 0];
                    const uint32_t lt = (1u << wl) - 1u;
                    uint32_t idx[SPL], my_cnt = 0, my_seg = 0, total = 0;
                    for (uint32_t d = 0; d <= a.g_mask; ++d) {
                        uint32_t c = 0;
#pragma unroll
                        for (int k = 0; k < SPL; ++k) {
                            const uint32_t bal = __ballot_sync(0xFFFFFFFFu, to[k] == d);
                            if (to[k] == d) idx[k] = c + __popc(bal & lt);
                            c += __popc(bal);
                        }
                        if ((uint32_t)wl == d) { my_cnt = c; my_seg = total; }
                        total += c;
                    }
                    uint32_t my_base = 0;
                    if (my_cnt) my_base = atomicAdd(a.view.ring_tail[me] + wl, my_cnt);
#pragma unroll
                    for (int k = 0; k < SPL; ++k) {
                        const uint32_t base = __shfl_sync(0xFFFFFFFFu, my_base, to[k] & 31u);
                        const uint32_t seg = __shfl_sync(0xFFFFFFFFu, my_seg, to[k] & 31u);
                        if (to[k] != 0xFFu) {
                            const uint32_t sp = seg + idx[k];
                            pack_message(pid[k], next[k], mu2[k], sg2[k], a.epoch, msgs.w[sp]);
                            msgs.at[sp] = base + idx[k];
                            msgs.to[sp] = (uint8_t)to[k];
                        }
                    }
                    __syncwarp();
                    for (uint32_t j = wl; j < total; j += 32u) {
                        const uint32_t at = msgs.at[j], r = msgs.to[j];
                        if (at < a.ring_cap) st_words_sys(a.view.ring[r] + ((size_t)me * a.ring_cap + at) * 4, msgs.w[j]);
                        else atomicOr(&a.ctrl[CTRL_SHARD], (unsigned long long)TS_SHARD_OVERFLOW);
                    }
                    __syncwarp();
                }
#pragma unroll
                for (int k = 0; k < SPL; ++k)
                    if (go && used[k] && ok && a.out.out_queue)
                        st_pair_stream(a.out.out_queue + ((size_t)om * P + lane + k * L) * 2, mu2[k], sg2[k], pol);
            } else {
#pragma unroll
                for (int k = 0; k < SPL; ++k) {
                    grp_mask[k] = __match_any_sync(0xFFFFFFFFu, to[k]);
                    place[k] = 0;
                    if (to[k] != 0xFFu && wl == __ffs(grp_mask[k]) - 1)
                        place[k] = atomicAdd(a.view.ring_tail[me] + to[k], (uint32_t)__popc(grp_mask[k]));
                }
#pragma unroll
                for (int k = 0; k < SPL; ++k) {
                    const int s = lane + k * L;
                    const uint32_t at = __shfl_sync(0xFFFFFFFFu, place[k], __ffs(grp_mask[k]) - 1) +
                                        __popc(grp_mask[k] & ((1u << wl) - 1u));
                    if (to[k] != 0xFFu) {
                        if (at < a.ring_cap)
                            st_message(a.view.ring[to[k]] + ((size_t)me * a.ring_cap + at) * 4, pid[k], next[k], mu2[k], sg2[k], a.epoch);
                        else atomicOr(&a.ctrl[CTRL_SHARD], (unsigned long long)TS_SHARD_OVERFLOW);
                    }
                    if (go && used[k] && ok && a.out.out_queue)
                        st_pair_stream(a.out.out_queue + ((size_t)om * P + s) * 2, mu2[k], sg2[k], pol);
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
    // The last rating block of this rank to finish tells every receiver how many messages this
    // sender left in its ring (the routers there stop when they have delivered that many).
    if (a.g_mask != 0u && !a.direct) {
        __syncthreads();
        if (threadIdx.x == 0) {
            uint32_t *tail = a.view.ring_tail[me];
            __threadfence();
            if (atomicAdd(tail + kTailDone, 1u) == blocks_of_rank - 1u) {
                __threadfence();
                tail[kTailDone] = 0;                               // for the next launch
                for (uint32_t g = 0; g <= a.g_mask; ++g) {
                    if (g == me) continue;
                    const unsigned long long f = ((unsigned long long)a.window_gen << 32) | atomicAdd(tail + g, 0u);
                    asm volatile("st.relaxed.sys.global.u64 [%0], %1;" ::"l"(a.view.ring_final[g] + me), "l"(f) : "memory");
                }
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
    // router blocks per rank: a sixteenth of the rank's grid (TSRATE_SHARD_ROUTERS overrides), at least one
    // warp per sender, and rating blocks must remain
    PushArgs a = args;
    const unsigned per_rank = emulate ? blocks >> a.g_shift : blocks;
    a.router_blocks = 0;
    if (a.g_mask != 0u && !a.direct) {
        const char *e = getenv("TSRATE_SHARD_ROUTERS");
        unsigned want = e ? (unsigned)atoi(e) : per_rank / 16u;
        const unsigned least = (a.g_mask + 7u) / 8u;
        want = want < least ? least : want;
        if (want + 1u > per_rank) { ts_set_error("ts_shard_rate: too few resident blocks for %u ranks", a.g_mask + 1u); return TS_E_UNSUPPORTED; }
        a.router_blocks = want;
    }
    dim3 grid(blocks), block(kBlock);
    void *kargs[] = {(void *)&a};
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
    if (view->n_shards > 1) {          // no message, no final count, no reservation
        if (!view->ring[view->rank] || !view->ring_final[view->rank] || !view->ring_tail[view->rank] || view->ring_cap < 1) {
            ts_set_error("ts_shard_reset: rings not mapped");
            return TS_E_ARG;
        }
        // a run's messages carry its epoch as marker: the rings are cleared only when the epochs start over
        if (view->ring_epoch <= 1)
            TS_CUDA(cudaMemsetAsync(view->ring[view->rank], 0, (size_t)view->n_shards * (size_t)view->ring_cap * 32, s));
        TS_CUDA(cudaMemsetAsync(view->ring_final[view->rank], 0, TS_SHARD_MAX * 8, s));
        TS_CUDA(cudaMemsetAsync(view->ring_tail[view->rank], 0, 3 * TS_SHARD_MAX * 4, s));
    }
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
        const bool mine = emulate_all || g == view->rank;
        if (!view->rows[g] || !view->table[g] || (mine && (!view->ids[g] || !view->meta[g] || !view->succ[g])) ||
            (G > 1 && (!view->ring[g] || !view->ring_final[g] || (mine && !view->ring_tail[g])))) {
            ts_set_error("ts_shard_rate: shard %d not mapped", g);
            return TS_E_ARG;
        }
    }
    if (G > 1 && (view->ring_cap < 1 || view->ring_cap > 0x7FFFFFFFll)) {
        ts_set_error("ts_shard_rate: ring_cap out of range");
        return TS_E_ARG;
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
    a.router_blocks = 0;
    a.ring_cap = (uint32_t)(G > 1 ? view->ring_cap : 0);
    a.window_gen = (uint32_t)(seq_base / TS_SHARD_BLOCK) + 1u;
    a.epoch = (uint32_t)view->ring_epoch;
    // Rows straight into the receivers' mailboxes while the peers' mailboxes together are small (scattered
    // stores over NVLink: 8.4 G stores/s into 320 MB of one peer, 5.7 G/s into 960 MB of three, 0.14 G/s into
    // 2.2 GB of seven); beyond 256 MB, messages in rings -- with the messages of a pass grouped by receiver
    // the rings are also the faster way into ONE peer's 320 MB (2 GPUs, 100M matches over 10M players: rating
    // 19.8 ms against 22.5 direct).  TSRATE_SHARD_DIRECT=0/1 overrides.
    const char *ed = getenv("TSRATE_SHARD_DIRECT");
    a.direct = ed ? (atoi(ed) != 0) : ((int64_t)(G - 1) * n_players * 32 <= (256ll << 20));
    // ring messages grouped by receiver in shared memory before they leave (TSRATE_SHARD_STAGE=0: as they come)
    const char *es = getenv("TSRATE_SHARD_STAGE");
    a.stage_msgs = es ? (atoi(es) != 0) : 1u;
    if (G > 1 && !a.direct && (a.epoch < 1u || a.epoch > 255u)) {
        ts_set_error("ts_shard_rate: ring_epoch must be 1..255 (ts_shard_reset hands it out)");
        return TS_E_ARG;
    }
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
