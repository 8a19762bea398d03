// GPU dependency scheduler for sm_100a.
//
// Replaces the sequential, chronological per-match loop of worker.py:176-192
// (state chains from one match to the next through the shared Player objects of
// the SQLAlchemy identity map) with an explicit schedule:
//
//   1. radix-sort the (player_id, slot index) pairs of all rated matches by
//      player id -- LSD, 8-bit digits, stable, so each player's slots stay in
//      match_seq order;
//   2. link each player's consecutive matches (predecessor match per slot);
//   3. rounds: round[m] = 1 + max over the slots of round[predecessor], evaluated
//      by dataflow along the links in one persistent kernel (k_levels; the
//      level-synchronous frontier sweep it replaced is kept behind an env switch);
//   4. radix-sort (round, match_seq) so every round lists its matches in
//      ascending match_seq, prefix-sum the level sizes into round_offsets, and
//      permute the match records into that order for the rating kernel.
//
// Gated matches (unsupported mode, AFK, invalid roster count) never touch a
// player (rater.py:83-85,102-106): they create no edges and land in bucket 0.
// All integer work; the result is bit-exact against the sequential schedule
// oracle by construction (longest-path levels).
#include <stdint.h>
#include <stdlib.h>

#include <algorithm>

#include "common.cuh"

namespace {

#ifndef TS_SORT_MIN_BLOCKS
#define TS_SORT_MIN_BLOCKS 5   // 48 registers, 5 x 42 KB shared: measured 2.75 ms link vs 2.94 (4) and 3.32 (3)
#endif
constexpr int kSortBlock = 256;
#ifndef TS_SORT_ITEMS
#define TS_SORT_ITEMS 16
#endif
constexpr int kSortItems = TS_SORT_ITEMS;
constexpr int kSortTile = kSortBlock * kSortItems;   // 4096 elements per block
constexpr int kWarps = kSortBlock / 32;

struct KeySrc {
    // pass 0 of the player sort reads match records directly: key = player id,
    // or n_players for slots that take no part (empty slot / gated match)
    const uint32_t *keys;
    const uint32_t *meta;
    ModeCols modes;
    uint32_t slots;
    uint32_t sentinel;
    int from_ids;
    unsigned long long slots_magic;   // floor(2^64 / slots) + 1: idx / slots == umul64hi(idx, magic) for idx < 2^32
    uint32_t rated_mask;   // bit m set: game mode m is rated (a register test instead of indexing modes.col)
};

__device__ __forceinline__ uint32_t load_key(const KeySrc &src, uint32_t idx)
{
    uint32_t k = __ldg(src.keys + idx);
    if (src.from_ids) {
        uint32_t m = __ldg(src.meta + (uint32_t)__umul64hi((unsigned long long)idx, src.slots_magic));
        const bool gated = ((src.rated_mask >> ((m >> TS_META_MODE_SHIFT) & 15u)) & 1u) == 0u ||
                           (m & (TS_META_AFK | TS_META_INVALID)) != 0u;
        if (k == TS_EMPTY_SLOT || gated) k = src.sentinel;
    }
    return k;
}

// per-block digit histogram, written digit-major: hist[d * n_blocks + block]
__global__ void __launch_bounds__(kSortBlock) k_radix_hist(KeySrc src, uint32_t n, int shift,
                                                           uint32_t *hist, uint32_t n_blocks)
{
    __shared__ uint32_t h[256];
    h[threadIdx.x] = 0;
    __syncthreads();
    const uint32_t base = blockIdx.x * kSortTile;
#pragma unroll 4
    for (int k = 0; k < kSortItems; ++k) {
        uint32_t idx = base + k * kSortBlock + threadIdx.x;
        if (idx < n) atomicAdd(&h[(load_key(src, idx) >> shift) & 255u], 1u);
    }
    __syncthreads();
    hist[threadIdx.x * n_blocks + blockIdx.x] = h[threadIdx.x];
}

// exclusive scan of each digit's row of block counts (one block per digit)
__global__ void __launch_bounds__(1024) k_radix_scan_rows(uint32_t *hist, uint32_t n_blocks, uint32_t *totals)
{
    __shared__ uint32_t warp_sums[32];
    __shared__ uint32_t carry;
    uint32_t *row = hist + (size_t)blockIdx.x * n_blocks;
    if (threadIdx.x == 0) carry = 0;
    __syncthreads();
    for (uint32_t base = 0; base < n_blocks; base += 1024) {
        uint32_t i = base + threadIdx.x;
        uint32_t x = i < n_blocks ? row[i] : 0u;
        uint32_t incl = x;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            uint32_t y = __shfl_up_sync(0xFFFFFFFFu, incl, o);
            if ((threadIdx.x & 31) >= o) incl += y;
        }
        if ((threadIdx.x & 31) == 31) warp_sums[threadIdx.x >> 5] = incl;
        __syncthreads();
        if (threadIdx.x < 32) {
            uint32_t w = warp_sums[threadIdx.x], wi = w;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                uint32_t y = __shfl_up_sync(0xFFFFFFFFu, wi, o);
                if (threadIdx.x >= o) wi += y;
            }
            warp_sums[threadIdx.x] = wi - w;
        }
        __syncthreads();
        uint32_t excl = carry + warp_sums[threadIdx.x >> 5] + incl - x;
        if (i < n_blocks) row[i] = excl;
        __syncthreads();
        if (threadIdx.x == 1023) carry = excl + x;
        __syncthreads();
    }
    if (threadIdx.x == 0) totals[blockIdx.x] = carry;
}

// exclusive scan of the 256 digit totals (one block)
__global__ void __launch_bounds__(256) k_radix_scan_totals(uint32_t *totals)
{
    __shared__ uint32_t s[256];
    s[threadIdx.x] = totals[threadIdx.x];
    __syncthreads();
    if (threadIdx.x == 0) {
        uint32_t run = 0;
        for (int d = 0; d < 256; ++d) { uint32_t c = s[d]; s[d] = run; run += c; }
    }
    __syncthreads();
    totals[threadIdx.x] = s[threadIdx.x];
}

// stable scatter: warp w of a block owns kSortItems*32 consecutive elements and
// walks them 32 at a time, so (step, lane) order is index order.  The tile is
// first put in digit order in shared memory, then written out with consecutive
// threads on consecutive addresses of each digit's run.
__global__ void __launch_bounds__(kSortBlock, 4)
k_radix_scatter(KeySrc src, const uint32_t *vals_in, int ident_vals, uint32_t n, int shift,
                const uint32_t *hist, const uint32_t *totals, uint32_t n_blocks,
                uint32_t *keys_out, uint32_t *vals_out)
{
    __shared__ uint32_t wcount[kWarps][256];
    __shared__ uint32_t gbase[256];       // global start of the digit's run for this block, minus its tile-local start
    __shared__ uint32_t warp_tot[kWarps];
    __shared__ uint32_t skey[kSortTile];
    __shared__ uint32_t sval[kSortTile];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    for (int i = threadIdx.x; i < kWarps * 256; i += kSortBlock) (&wcount[0][0])[i] = 0;
    __syncthreads();
    const uint32_t tile_base = blockIdx.x * kSortTile;
    const uint32_t wbase = tile_base + warp * (kSortItems * 32);
    uint32_t key[kSortItems], off[kSortItems];
    // all loads first (16 independent requests per thread in flight) ...
#pragma unroll
    for (int k = 0; k < kSortItems; ++k) {
        uint32_t idx = wbase + k * 32 + lane;
        key[k] = idx < n ? load_key(src, idx) : 0u;
    }
    // ... then the ranks: lanes with equal digits find each other (ballots, below);
    // the lowest such lane bumps the warp's counter of that digit with a shared-
    // memory atomic (same-address atomics of one warp retire in program order, so
    // later steps rank behind earlier ones: stable)
#pragma unroll
    for (int k = 0; k < kSortItems; ++k) {
        uint32_t idx = wbase + k * 32 + lane;
        bool ok = idx < n;
        const uint32_t d = (key[k] >> shift) & 255u;
        // lanes holding the same digit: one ballot per digit bit (match.any walks the
        // distinct values of the warp one by one and is several times slower here)
        uint32_t peers = __ballot_sync(0xFFFFFFFFu, ok);
        peers = ok ? peers : ~peers;
#pragma unroll
        for (int b = 0; b < 8; ++b) {
            const bool bit = (d >> b) & 1u;
            const uint32_t bal = __ballot_sync(0xFFFFFFFFu, bit);
            peers &= bit ? bal : ~bal;
        }
        int leader = __ffs(peers) - 1;
        uint32_t prev = 0;
        if (ok && lane == leader) prev = atomicAdd(&wcount[warp][d], (uint32_t)__popc(peers));
        off[k] = __shfl_sync(0xFFFFFFFFu, prev, leader) + __popc(peers & ((1u << lane) - 1u));
    }
    __syncthreads();
    {
        // thread d: per-warp starts inside digit d's run, then the run's tile-local start
        const uint32_t d = threadIdx.x;
        uint32_t run = 0;
#pragma unroll
        for (int w = 0; w < kWarps; ++w) {
            uint32_t c = wcount[w][d];
            wcount[w][d] = run;
            run += c;
        }
        uint32_t incl = run;   // exclusive scan of the 256 digit totals across the block
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            uint32_t y = __shfl_up_sync(0xFFFFFFFFu, incl, o);
            if (lane >= o) incl += y;
        }
        if (lane == 31) warp_tot[warp] = incl;
        __syncthreads();
        uint32_t before = 0;
#pragma unroll
        for (int w = 0; w < kWarps; ++w) before += (w < warp) ? warp_tot[w] : 0u;
        const uint32_t dstart = before + incl - run;
        gbase[d] = __ldg(totals + d) + __ldg(hist + (size_t)d * n_blocks + blockIdx.x) - dstart;
#pragma unroll
        for (int w = 0; w < kWarps; ++w) wcount[w][d] += dstart;
    }
    __syncthreads();
#pragma unroll
    for (int k = 0; k < kSortItems; ++k) {
        uint32_t idx = wbase + k * 32 + lane;
        if (idx < n) {
            uint32_t pos = wcount[warp][(key[k] >> shift) & 255u] + off[k];
            skey[pos] = key[k];
            sval[pos] = ident_vals ? idx : __ldg(vals_in + idx);
        }
    }
    __syncthreads();
    const uint32_t cnt = min((uint32_t)kSortTile, n - tile_base);
    for (uint32_t i = threadIdx.x; i < cnt; i += kSortBlock) {
        uint32_t k = skey[i];
        uint32_t pos = gbase[(k >> shift) & 255u] + i;
        keys_out[pos] = k;
        vals_out[pos] = sval[i];
    }
}

// ---------------------------------------------------------------------------
// Single-sweep form of a radix pass (decoupled look-back, in the manner of
// Merrill & Garland's scan and Adinets & Merrill's "onesweep" sort): the digit
// totals of all passes are counted once up front (keys do not change from pass to
// pass), and each scatter block publishes its per-digit counts in a status row and
// looks back over the rows of earlier tiles for its offsets.  No per-pass histogram
// kernel, no scan kernels, keys are read once per pass.  Tile numbers are handed
// out by an atomic ticket so that every tile a block waits for has already started.
// Status word: bit 31 = inclusive prefix available, bit 30 = tile count available,
// low 30 bits the value (so this form needs n < 2^30).
// ---------------------------------------------------------------------------
constexpr uint32_t kStPrefix = 0x80000000u, kStCount = 0x40000000u, kStMask = 0x3FFFFFFFu;
constexpr int kMaxPasses = 4;
#ifndef TS_SORT_LOOKBACK
#define TS_SORT_LOOKBACK 1
#endif
constexpr int kLookBack = TS_SORT_LOOKBACK;
constexpr int kTotalsWords = kMaxPasses * 256 + 64;    // digit bases of every pass + the ticket

__global__ void __launch_bounds__(kSortBlock)
k_radix_hist_all(KeySrc src, uint32_t n_in, const uint32_t *n_dev, int passes, uint32_t *ghist)
{
    const uint32_t n = n_dev ? __ldg(n_dev) : n_in;
    __shared__ uint32_t h[kMaxPasses][256];
    for (int i = threadIdx.x; i < kMaxPasses * 256; i += kSortBlock) (&h[0][0])[i] = 0;
    __syncthreads();
    for (uint32_t idx = blockIdx.x * kSortBlock + threadIdx.x; idx < n; idx += gridDim.x * kSortBlock) {
        const uint32_t k = load_key(src, idx);
        for (int p = 0; p < passes; ++p) atomicAdd(&h[p][(k >> (8 * p)) & 255u], 1u);
    }
    __syncthreads();
    for (int i = threadIdx.x; i < passes * 256; i += kSortBlock) {
        const uint32_t c = (&h[0][0])[i];
        if (c) atomicAdd(&ghist[i], c);
    }
}

// The same totals for pass 0's source, the match records themselves: one thread per match reads
// the meta word once and the ids two at a time -- no division by the slot count, no gather.
__global__ void __launch_bounds__(kSortBlock)
k_radix_hist_matches(KeySrc src, uint32_t n_matches, int passes, uint32_t *ghist)
{
    __shared__ uint32_t h[kMaxPasses][256];
    for (int i = threadIdx.x; i < kMaxPasses * 256; i += kSortBlock) (&h[0][0])[i] = 0;
    __syncthreads();
    const uint32_t pairs = src.slots / 2;
    for (uint32_t m = blockIdx.x * kSortBlock + threadIdx.x; m < n_matches; m += gridDim.x * kSortBlock) {
        const uint32_t meta = __ldg(src.meta + m);
        const bool gated = ((src.rated_mask >> ((meta >> TS_META_MODE_SHIFT) & 15u)) & 1u) == 0u ||
                           (meta & (TS_META_AFK | TS_META_INVALID)) != 0u;
        const uint2 *row = reinterpret_cast<const uint2 *>(src.keys + (size_t)m * src.slots);
        for (uint32_t j = 0; j < pairs; ++j) {
            const uint2 two = __ldg(row + j);
#pragma unroll
            for (int e = 0; e < 2; ++e) {
                uint32_t k = e ? two.y : two.x;
                if (k == TS_EMPTY_SLOT || gated) k = src.sentinel;
                for (int p = 0; p < passes; ++p) atomicAdd(&h[p][(k >> (8 * p)) & 255u], 1u);
            }
        }
    }
    __syncthreads();
    for (int i = threadIdx.x; i < passes * 256; i += kSortBlock) {
        const uint32_t c = (&h[0][0])[i];
        if (c) atomicAdd(&ghist[i], c);
    }
}

// exclusive scan of each pass's 256 digit totals (one block, thread = digit)
__global__ void __launch_bounds__(256) k_radix_scan_all(uint32_t *ghist, int passes)
{
    __shared__ uint32_t s[256];
    for (int p = 0; p < passes; ++p) {
        s[threadIdx.x] = ghist[p * 256 + threadIdx.x];
        __syncthreads();
        if (threadIdx.x == 0) {
            uint32_t run = 0;
            for (int d = 0; d < 256; ++d) { uint32_t c = s[d]; s[d] = run; run += c; }
        }
        __syncthreads();
        ghist[p * 256 + threadIdx.x] = s[threadIdx.x];
        __syncthreads();
    }
}

__device__ __forceinline__ uint32_t ld_relaxed_u32(const uint32_t *p)
{
    uint32_t v;
    asm volatile("ld.relaxed.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_relaxed_u32(uint32_t *p, uint32_t v)
{
    asm volatile("st.relaxed.gpu.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}

#ifndef TS_SORT_LATE_LOOKBACK
#define TS_SORT_LATE_LOOKBACK 0
#endif
#ifndef TS_SORT_SPIN_NS
#define TS_SORT_SPIN_NS 40
#endif

// Decoupled look-back of thread d = digit d: sum the counts of the tiles before `tile` back to the
// nearest published prefix, publish this tile's prefix, and leave in gbase[d] where this tile's
// run of digit d starts in the output, minus its tile-local start.
// kLookBack > 1 loads the statuses of several predecessors together (independent L2 round trips)
// before consuming them in order; measured slower (config 2 link 2.59 -> 2.70 ms at 4, 2.76 at 8):
// the loop mostly waits for the tile just before to publish, it does not walk far.
__device__ __forceinline__ void lookback(uint32_t *status, uint32_t tile, uint32_t d, uint32_t run, uint32_t dstart,
                                         const uint32_t *digit_base, uint32_t *gbase)
{
    uint32_t excl = 0;
    for (uint32_t j = tile; j > 0;) {
        uint32_t v[kLookBack];
#pragma unroll
        for (int q = 0; q < kLookBack; ++q)
            v[q] = j > (uint32_t)q ? ld_relaxed_u32(status + (size_t)(j - 1 - q) * 256 + d) : kStPrefix;   // before tile 0: prefix 0
        bool found = false, pending = false;
#pragma unroll
        for (int q = 0; q < kLookBack; ++q) {
            if (found || pending) continue;
            if (v[q] & kStPrefix) { excl += v[q] & kStMask; found = true; }
            else if (v[q] & kStCount) { excl += v[q] & kStMask; --j; }
            else pending = true;
        }
        if (found) break;
        if (pending) __nanosleep(TS_SORT_SPIN_NS);  // that tile has not published yet: do not hog the issue slots
    }
    if (tile > 0) st_relaxed_u32(status + (size_t)tile * 256 + d, (excl + run) | kStPrefix);
    gbase[d] = __ldg(digit_base + d) + excl - dstart;
}

// n_dev (optional): the number of entries lives on the device (the owned slots of a sharded ladder)
__global__ void __launch_bounds__(kSortBlock, TS_SORT_MIN_BLOCKS)
k_radix_onesweep(KeySrc src, const uint32_t *vals_in, int ident_vals, uint32_t n_in, const uint32_t *n_dev, int shift,
                 const uint32_t *digit_base, uint32_t *status, uint32_t *ticket, uint32_t *keys_out,
                 uint32_t *vals_out, const uint32_t *key_bound)
{
    // key_bound (optional, device): the largest key.  A pass whose digit is zero for every key
    // would only copy: it is skipped, and the consumers find the result one buffer earlier
    // (result_buffer below) -- the pass count is decided on the device, the host never waits.
    if (key_bound && shift > 0 && (__ldg(key_bound) >> shift) == 0u) return;
    const uint32_t n = n_dev ? __ldg(n_dev) : n_in;
    __shared__ uint32_t wcount[kWarps][256];
    __shared__ uint32_t gbase[256];
    __shared__ uint32_t warp_tot[kWarps];
    __shared__ uint32_t skey[kSortTile];
    __shared__ uint32_t sval[kSortTile];
    __shared__ uint32_t s_tile;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    if (threadIdx.x == 0) s_tile = atomicAdd(ticket, 1u);
    for (int i = threadIdx.x; i < kWarps * 256; i += kSortBlock) (&wcount[0][0])[i] = 0;
    __syncthreads();
    const uint32_t tile = s_tile;
    const uint32_t tile_base = tile * kSortTile;
    if (tile_base >= n) return;            // beyond the device-side count: no earlier tile waits for this one
    const uint32_t wbase = tile_base + warp * (kSortItems * 32);
    uint32_t key[kSortItems], off[kSortItems];
#pragma unroll
    for (int k = 0; k < kSortItems; ++k) {
        uint32_t idx = wbase + k * 32 + lane;
        key[k] = idx < n ? load_key(src, idx) : 0u;
    }
#pragma unroll
    for (int k = 0; k < kSortItems; ++k) {
        uint32_t idx = wbase + k * 32 + lane;
        bool ok = idx < n;
        const uint32_t d = (key[k] >> shift) & 255u;
        uint32_t peers = __ballot_sync(0xFFFFFFFFu, ok);
        peers = ok ? peers : ~peers;
#pragma unroll
        for (int b = 0; b < 8; ++b) {
            const bool bit = (d >> b) & 1u;
            const uint32_t bal = __ballot_sync(0xFFFFFFFFu, bit);
            peers &= bit ? bal : ~bal;
        }
        int leader = __ffs(peers) - 1;
        uint32_t prev = 0;
        if (ok && lane == leader) prev = atomicAdd(&wcount[warp][d], (uint32_t)__popc(peers));
        off[k] = __shfl_sync(0xFFFFFFFFu, prev, leader) + __popc(peers & ((1u << lane) - 1u));
    }
    __syncthreads();
    uint32_t my_run, my_dstart;
    {
        const uint32_t d = threadIdx.x;
        uint32_t run = 0;
#pragma unroll
        for (int w = 0; w < kWarps; ++w) {
            uint32_t c = wcount[w][d];
            wcount[w][d] = run;
            run += c;
        }
        // publish this tile's count of digit d; the tiles before it are looked at later
        uint32_t *row = status + (size_t)tile * 256;
        st_relaxed_u32(row + d, run | (tile == 0 ? kStPrefix : kStCount));
        uint32_t incl = run;   // exclusive scan of the 256 digit totals across the block (tile-local starts)
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            uint32_t y = __shfl_up_sync(0xFFFFFFFFu, incl, o);
            if (lane >= o) incl += y;
        }
        if (lane == 31) warp_tot[warp] = incl;
        __syncthreads();
        uint32_t before = 0;
#pragma unroll
        for (int w = 0; w < kWarps; ++w) before += (w < warp) ? warp_tot[w] : 0u;
        const uint32_t dstart = before + incl - run;
#pragma unroll
        for (int w = 0; w < kWarps; ++w) wcount[w][d] += dstart;
        my_run = run;
        my_dstart = dstart;
    }
    __syncthreads();
    // Look-back first, then the tile goes into digit order in shared memory.  The other order
    // (TS_SORT_LATE_LOOKBACK=1: reorder while the tiles before publish) measured slower -- config 2 link
    // 2.58 -> 2.82 ms: what the tiles behind wait for is this tile's prefix, and that came later.  The
    // back-off of the spin (TS_SORT_SPIN_NS 40 / 100 / 200) makes no difference.
#if !TS_SORT_LATE_LOOKBACK
    lookback(status, tile, threadIdx.x, my_run, my_dstart, digit_base, gbase);
    __syncthreads();
#endif
#pragma unroll
    for (int k = 0; k < kSortItems; ++k) {
        uint32_t idx = wbase + k * 32 + lane;
        if (idx < n) {
            uint32_t pos = wcount[warp][(key[k] >> shift) & 255u] + off[k];
            skey[pos] = key[k];
            sval[pos] = ident_vals ? idx : __ldg(vals_in + idx);
        }
    }
#if TS_SORT_LATE_LOOKBACK
    lookback(status, tile, threadIdx.x, my_run, my_dstart, digit_base, gbase);
#endif
    __syncthreads();
    const uint32_t cnt = min((uint32_t)kSortTile, n - tile_base);
    for (uint32_t i = threadIdx.x; i < cnt; i += kSortBlock) {
        uint32_t k = skey[i];
        uint32_t pos = gbase[(k >> shift) & 255u] + i;
        keys_out[pos] = k;
        vals_out[pos] = sval[i];
    }
}

// ---- chain linking --------------------------------------------------------
// Sorted pairs (key = player, val = slot index i = m*P + s).  Runs of equal key
// list one player's slots in match order.  The first entry of a (player, match)
// run is its primary: it records whether an earlier match of the same player
// exists (bit 31) and carries the link to the player's next match (bits 0..30).
constexpr uint32_t kLinkNone = 0x7FFFFFFFu;
constexpr uint32_t kLinkHasPred = 0x80000000u;

__global__ void __launch_bounds__(256)
k_link(const uint32_t *keys, const uint32_t *vals, uint32_t n, uint32_t slots, uint32_t sentinel, uint32_t *succ)
{
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    uint32_t k = keys[i], v = vals[i];
    if (k >= sentinel) { succ[v] = kLinkNone; return; }
    uint32_t m = v / slots;
    bool has_prev = i > 0 && keys[i - 1] == k;
    if (has_prev && vals[i - 1] / slots == m) { succ[v] = kLinkNone; return; }   // duplicate of a player inside one match
    uint32_t nx = kLinkNone;
    for (uint32_t j = i + 1; j < n && keys[j] == k; ++j) {
        uint32_t mj = vals[j] / slots;
        if (mj != m) { nx = mj; break; }
    }
    succ[v] = nx | (has_prev ? kLinkHasPred : 0u);
}

// pending[m] = number of slots with a predecessor; matches with none start the sweep
__global__ void __launch_bounds__(256)
k_init_frontier(const uint32_t *meta, ModeCols modes, const uint32_t *succ, uint32_t slots, uint32_t n,
                uint32_t *pending, uint32_t *queue, unsigned long long *ctrl)
{
    uint32_t m = blockIdx.x * blockDim.x + threadIdx.x;
    bool ready = false;
    if (m < n) {
        uint32_t cnt = 0;
        for (uint32_t s = 0; s < slots; ++s) cnt += succ[(size_t)m * slots + s] >> 31;
        pending[m] = cnt;
        ready = cnt == 0u && !meta_is_gated(__ldg(meta + m), modes.col);
    }
    // warp-aggregated append
    uint32_t mask = __ballot_sync(0xFFFFFFFFu, ready);
    if (mask == 0u) return;
    int lane = threadIdx.x & 31, leader = __ffs(mask) - 1;
    unsigned long long base = 0;
    if (lane == leader) base = atomicAdd(&ctrl[CTRL_TAIL], (unsigned long long)__popc(mask));
    base = __shfl_sync(0xFFFFFFFFu, base, leader);
    if (ready) queue[base + __popc(mask & ((1u << lane) - 1u))] = m;
}

// Level-synchronous frontier sweep.  queue holds level after level back to back.
// Level L+1's entries are claimed through a per-level word whose low half counts
// the appended matches and whose high half counts the blocks that have finished
// level L; a block polls that one word until the whole grid has arrived and
// reads the size of the next level from the same load.  Three words rotate.
__device__ __forceinline__ void prefetch_l2(const void *p) { asm volatile("prefetch.global.L2 [%0];" ::"l"(p)); }

__global__ void __launch_bounds__(256)
k_frontier(uint32_t *queue, const uint32_t *succ, uint32_t *pending, uint32_t *round, uint32_t slots,
           uint32_t n_matches, uint32_t *level_sizes, uint32_t offsets_cap, ts_sched_info *info,
           unsigned long long *ctrl)
{
    __shared__ uint32_t s_cnt;
    uint32_t start = 0;
    // the host copied the tail after k_init_frontier into the snapshot word
    uint32_t end = (uint32_t)ld_acquire_u64(&ctrl[CTRL_TAIL + 1]);
    uint32_t level = 1;
    const int lane = threadIdx.x & 31;
    while (start < end) {
        unsigned long long *word = &ctrl[CTRL_LEVEL + level % 3];      // collects level + 1
        const uint32_t work = (end - start) * slots;
        const uint32_t stride = gridDim.x * blockDim.x;
        // whole warps iterate together so the appends can be warp-aggregated
        for (uint32_t i0 = blockIdx.x * blockDim.x + (threadIdx.x & ~31u); i0 < work; i0 += stride) {
            const uint32_t i = i0 + lane;
            bool push = false;
            uint32_t nx = kLinkNone;
            if (i < work) {
                uint32_t m = __ldcg(queue + start + i / slots);
                uint32_t s = i % slots;
                if (s == 0) round[m] = level;
                nx = __ldg(succ + (size_t)m * slots + s) & kLinkNone;
                push = nx != kLinkNone && atomicSub(&pending[nx], 1u) == 1u;
            }
            uint32_t mask = __ballot_sync(0xFFFFFFFFu, push);
            if (mask) {
                int leader = __ffs(mask) - 1;
                unsigned long long base = 0;
                if (lane == leader) base = atomicAdd(word, (unsigned long long)__popc(mask));
                base = __shfl_sync(0xFFFFFFFFu, base, leader);
                if (push) {
                    queue[end + (uint32_t)base + __popc(mask & ((1u << lane) - 1u))] = nx;
                    // next level reads this match's links right after the barrier
                    prefetch_l2(succ + (size_t)nx * slots);
                    prefetch_l2(succ + (size_t)nx * slots + slots - 1);
                }
            }
        }
        __syncthreads();
        if (threadIdx.x == 0) {
            __threadfence();
            red_add_u64(word, 1ull << 32);
            unsigned long long w;
            while (((w = ld_relaxed_u64(word)) >> 32) < gridDim.x) {
            }
            __threadfence();
            s_cnt = (uint32_t)w;
            if (blockIdx.x == 0) {
                if (level + 2 <= offsets_cap) level_sizes[level] = end - start;
                // the word of two levels ago is idle: everyone polled it before arriving here
                ctrl[CTRL_LEVEL + (level + 2) % 3] = 0ull;
            }
        }
        __syncthreads();
        start = end;
        end = start + s_cnt;
        level += 1;
    }
    if (blockIdx.x == 0 && threadIdx.x == 0) {
        info->n_rounds = level - 1;
        info->n_gated = n_matches - end;
        info->overflow = (level - 1 + 2 > offsets_cap) ? 1u : 0u;
        info->reserved = 0;
        if (offsets_cap >= 1) level_sizes[0] = n_matches - end;
    }
}

// in-place exclusive scan of bucket sizes -> round_offsets (single block; the
// number of rounds is small next to the number of matches)
__global__ void __launch_bounds__(1024) k_offsets(uint32_t *sizes, const ts_sched_info *info, uint32_t cap)
{
    __shared__ uint32_t warp_sums[32];
    __shared__ uint32_t carry;
    uint32_t n = info->n_rounds + 1;           // buckets 0..n_rounds
    if (n + 1 > cap) n = cap - 1;
    if (threadIdx.x == 0) carry = 0;
    __syncthreads();
    for (uint32_t base = 0; base < n; base += 1024) {
        uint32_t i = base + threadIdx.x;
        uint32_t x = i < n ? sizes[i] : 0u, incl = x;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            uint32_t y = __shfl_up_sync(0xFFFFFFFFu, incl, o);
            if ((threadIdx.x & 31) >= o) incl += y;
        }
        if ((threadIdx.x & 31) == 31) warp_sums[threadIdx.x >> 5] = incl;
        __syncthreads();
        if (threadIdx.x < 32) {
            uint32_t w = warp_sums[threadIdx.x], wi = w;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                uint32_t y = __shfl_up_sync(0xFFFFFFFFu, wi, o);
                if (threadIdx.x >= o) wi += y;
            }
            warp_sums[threadIdx.x] = wi - w;
        }
        __syncthreads();
        uint32_t excl = carry + warp_sums[threadIdx.x >> 5] + incl - x;
        if (i < n) sizes[i] = excl;
        __syncthreads();
        if (threadIdx.x == 1023) carry = excl + x;
        __syncthreads();
    }
    if (threadIdx.x == 0) sizes[n] = carry;
}

__global__ void __launch_bounds__(256)
k_permute(const uint32_t *ids, const uint32_t *meta, const uint32_t *order, uint32_t n, uint32_t slots,
          uint32_t *ids_sorted, uint32_t *meta_sorted)
{
    // one thread per slot: contiguous writes, record-granular gathers
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= (size_t)n * slots) return;
    uint32_t j = (uint32_t)(i / slots), s = (uint32_t)(i % slots);
    uint32_t m = __ldg(order + j);
    ids_sorted[i] = __ldg(ids + (size_t)m * slots + s);
    if (s == 0) meta_sorted[j] = __ldg(meta + m);
}

// Predecessor links for the streaming (dataflow) rating kernel: pred[v] = index of
// the previous rated match of the player in slot v, or TS_NONE.  The scatter by
// slot index is random; it is done in passes over destination ranges small
// enough to stay in L2, so every 32-byte sector of pred is written back to HBM
// once, complete, instead of being read-modified-written eight times.
__global__ void __launch_bounds__(256)
k_pred(const uint32_t *keys, const uint32_t *vals, uint32_t n_in, const uint32_t *n_dev, unsigned long long slots_magic,
       uint32_t sentinel, uint32_t v_lo, uint32_t v_hi, uint32_t seq_base, const uint32_t *last, uint32_t *pred)
{
    const uint32_t n = n_dev ? __ldg(n_dev) : n_in;
    // four consecutive sorted entries per thread: two 16-byte streaming loads plus the
    // entry just before them, all issued up front
    const uint32_t i0 = (blockIdx.x * blockDim.x + threadIdx.x) * 4u;
    if (i0 >= n) return;
    uint32_t k[5], v[5];
    if (i0 + 4 <= n) {
        const uint4 kk = __ldcs(reinterpret_cast<const uint4 *>(keys + i0));
        const uint4 vv = __ldcs(reinterpret_cast<const uint4 *>(vals + i0));
        k[1] = kk.x; k[2] = kk.y; k[3] = kk.z; k[4] = kk.w;
        v[1] = vv.x; v[2] = vv.y; v[3] = vv.z; v[4] = vv.w;
    } else {
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const bool ok = i0 + j < n;
            k[j + 1] = ok ? __ldcs(keys + i0 + j) : 0xFFFFFFFFu;
            v[j + 1] = ok ? __ldcs(vals + i0 + j) : 0u;
        }
    }
    k[0] = i0 > 0 ? __ldcs(keys + i0 - 1) : ~k[1];
    v[0] = i0 > 0 ? __ldcs(vals + i0 - 1) : 0u;
#pragma unroll
    for (int j = 1; j <= 4; ++j) {
        const uint32_t i = i0 + j - 1;
        if (i >= n || v[j] < v_lo || v[j] >= v_hi || k[j] >= sentinel) continue;   // other pass / empty slot / gated match
        const uint32_t m = (uint32_t)__umul64hi((unsigned long long)v[j], slots_magic);
        uint32_t pm = TS_NONE;      // local index of the previous match in this window
        if (k[j - 1] == k[j]) {
            pm = (uint32_t)__umul64hi((unsigned long long)v[j - 1], slots_magic);
            if (pm == m) {          // the player is listed twice in this match: walk back to a different match
                pm = TS_NONE;
                for (uint32_t q = i - 1; q > 0 && __ldcs(keys + q - 1) == k[j]; --q) {
                    uint32_t mq = (uint32_t)__umul64hi((unsigned long long)__ldcs(vals + q - 1), slots_magic);
                    if (mq != m) { pm = mq; break; }
                }
            }
        }
        // stream-wide match numbers; a player's first match of this window continues the
        // chain at the tail carried over from earlier windows
        pm = pm != TS_NONE ? seq_base + pm : (last ? __ldg(last + k[j]) : TS_NONE);
        pred[v[j]] = pm;
    }
}

// after all pred passes: the last slot of every player's run becomes the carried tail
__global__ void __launch_bounds__(256)
k_last_update(const uint32_t *keys, const uint32_t *vals, uint32_t n_in, const uint32_t *n_dev,
              unsigned long long slots_magic, uint32_t sentinel, uint32_t seq_base, uint32_t *last)
{
    const uint32_t n = n_dev ? __ldg(n_dev) : n_in;
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const uint32_t k = __ldcs(keys + i);
    if (k >= sentinel) return;
    if (i + 1 < n && __ldcs(keys + i + 1) == k) return;
    last[k] = seq_base + (uint32_t)__umul64hi((unsigned long long)__ldcs(vals + i), slots_magic);
}

// ---- rounds by dataflow -----------------------------------------------------
// round[m] = 1 + max over the match's players of the round of that player's previous rated
// match (0 for gated matches) is a recurrence along the links, so it is evaluated the way the
// streaming rating kernel evaluates ratings: matches in stream order, one lane per match, a
// lane polls round[] of its predecessors until none is 0 ("not known yet") and publishes its
// own.  A round word validates itself.  No frontier queue, no atomics per edge, no grid-wide
// step per round: a dependency hop costs about one L2 round trip and a half (config 2, 888
// rounds: 4.3 ms of level-synchronous sweep before).
// The whole grid must be resident (cooperative launch); a warp takes its tiles of 32 matches
// in increasing order, so the oldest unresolved match never waits on work that has not started.
template <int P>     // P = slots (compile time for 6 and 10), 0 = run-time slot count
__global__ void __launch_bounds__(256)
k_levels(const uint32_t *pred, const uint32_t *meta, uint32_t rated_mask, uint32_t n, uint32_t slots_rt,
         uint32_t *round, ts_sched_info *info)
{
    const uint32_t slots = P ? (uint32_t)P : slots_rt;
    const int lane = threadIdx.x & 31;
    const uint32_t n_warps = gridDim.x * (blockDim.x >> 5);
    const uint32_t warp = (threadIdx.x >> 5) * gridDim.x + blockIdx.x;   // neighbouring tiles on different SMs
    const uint32_t n_tiles = (n + 31u) / 32u;
    uint32_t deepest = 0, gated_seen = 0;
    for (uint32_t tile = warp; tile < n_tiles; tile += n_warps) {
        const uint32_t m = tile * 32u + lane;
        bool live = false;
        if (m < n) {
            const uint32_t mt = __ldg(meta + m);
            live = ((rated_mask >> ((mt >> TS_META_MODE_SHIFT) & 15u)) & 1u) != 0u &&
                   (mt & (TS_META_AFK | TS_META_INVALID)) == 0u;
        }
        gated_seen += __popc(__ballot_sync(0xFFFFFFFFu, m < n && !live));
        // the predecessors still unknown, as a bit per slot
        uint32_t todo = 0, best = 0;
        const uint32_t *links = pred + (size_t)m * slots;
        if (live) {
#pragma unroll
            for (uint32_t sl = 0; sl < (P ? (uint32_t)P : 16u); ++sl)
                if (sl < slots && __ldg(links + sl) != TS_NONE) todo |= 1u << sl;
        }
        bool done = !live;
        while (!__all_sync(0xFFFFFFFFu, done)) {
            if (!done) {
                uint32_t lv[P ? P : 16];
#pragma unroll
                for (uint32_t sl = 0; sl < (P ? (uint32_t)P : 16u); ++sl)
                    lv[sl] = ((todo >> sl) & 1u) ? ld_relaxed_u32(round + __ldg(links + sl)) : 0u;
#pragma unroll
                for (uint32_t sl = 0; sl < (P ? (uint32_t)P : 16u); ++sl)
                    if (lv[sl] != 0u) { best = max(best, lv[sl]); todo &= ~(1u << sl); }
                if (todo == 0u) {
                    st_relaxed_u32(round + m, best + 1u);
                    deepest = max(deepest, best + 1u);
                    done = true;
                }
            }
        }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) deepest = max(deepest, __shfl_xor_sync(0xFFFFFFFFu, deepest, o));
    if (lane == 0) {
        if (deepest) atomicMax(&info->n_rounds, deepest);
        if (gated_seen) atomicAdd(&info->n_gated, gated_seen);
    }
}

// which of the two ping-pong buffers holds a sort whose passes beyond the largest key were skipped
__device__ __forceinline__ int result_buffer(uint32_t key_bound)
{
    int passes = 1;
    while (passes < 4 && (key_bound >> (8 * passes)) != 0u) ++passes;
    return (passes - 1) & 1;
}

// round_offsets from the rounds in sorted order: bucket r starts where the first key >= r stands;
// order = the match indices in that order.  The number of rounds is read on the device.
__global__ void __launch_bounds__(256)
k_round_bounds(const uint32_t *keys0, const uint32_t *keys1, const uint32_t *vals0, const uint32_t *vals1, uint32_t n,
               ts_sched_info *info, uint32_t cap, uint32_t *round_offsets, uint32_t *order)
{
    const uint32_t n_rounds = info->n_rounds;
    const int buf = result_buffer(n_rounds);
    const uint32_t *sorted_round = buf ? keys1 : keys0;
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i == 0 && (uint64_t)n_rounds + 2 > (uint64_t)cap) info->overflow = 1;
    if (i > n) return;
    if (i < n) order[i] = __ldg((buf ? vals1 : vals0) + i);
    const uint32_t lo = i == 0 ? 0u : __ldg(sorted_round + i - 1) + 1u;           // first bucket starting at i
    const uint32_t hi = i == n ? n_rounds + 1u : __ldg(sorted_round + i);         // last one
    for (uint32_t r = lo; r <= hi; ++r)
        if (r < cap) round_offsets[r] = i;
}

struct Workspace {
    uint32_t *keys[2];
    uint32_t *vals[2];
    uint32_t *hist;
    uint32_t *totals;
    uint32_t *succ;
    uint32_t *pending;
    uint32_t *queue;
    size_t bytes;
};

size_t align_up(size_t x) { return (x + 255) & ~(size_t)255; }

Workspace carve(char *base, int64_t n_matches, int32_t slots)
{
    Workspace w;
    size_t np = (size_t)n_matches * slots;
    size_t n_blocks = (np + kSortTile - 1) / kSortTile;
    size_t off = 0;
    auto take = [&](size_t bytes) { char *p = base ? base + off : nullptr; off += align_up(bytes); return (uint32_t *)p; };
    for (int i = 0; i < 2; ++i) { w.keys[i] = take(np * 4); w.vals[i] = take(np * 4); }
    w.hist = take((n_blocks * 256 + 256) * 4);
    w.totals = take(kTotalsWords * 4);
    w.succ = take(np * 4);
    w.pending = take((size_t)n_matches * 4);
    w.queue = take((size_t)n_matches * 4);
    w.bytes = off;
    return w;
}

// sort buffers only (ts_link_pred, ts_shard_link): two (key, val) buffers of np entries, the
// tile status rows and the digit totals
Workspace carve_link(char *base, uint32_t np)
{
    Workspace w;
    size_t off = 0, n_blocks = ((size_t)np + kSortTile - 1) / kSortTile;
    auto take = [&](size_t bytes) { char *p = base + off; off += align_up(bytes); return (uint32_t *)p; };
    for (int i = 0; i < 2; ++i) { w.keys[i] = take((size_t)np * 4); w.vals[i] = take((size_t)np * 4); }
    w.hist = take((n_blocks * 256 + 256) * 4);
    w.totals = take(kTotalsWords * 4);
    w.succ = w.pending = w.queue = nullptr;
    w.bytes = off;
    return w;
}

int frontier_blocks_per_sm_cap()
{
    static int cap = -1;
    if (cap < 0) {
        const char *e = getenv("TSRATE_FRONTIER_BLOCKS_PER_SM");
        cap = e ? atoi(e) : 2;
        if (cap < 1) cap = 1;
    }
    return cap;
}

uint32_t pred_range_entries()
{
    static uint32_t mi = 0;
    if (mi == 0) {
        const char *e = getenv("TSRATE_PRED_RANGE_MI");   // Mi entries (4 bytes each) of pred per pass
        int v = e ? atoi(e) : 16;
        mi = (uint32_t)(v < 1 ? 1 : (v > 512 ? 512 : v));
    }
    return mi << 20;
}

int bits_for(uint64_t max_value)
{
    int b = 1;
    while (b < 32 && (max_value >> b) != 0) ++b;
    return b;
}

// LSD radix sort of n (key, val) pairs on the low `bits` bits.  Pass 0 reads
// `src` (with identity values); results ping-pong between the two buffers.
// Returns the index of the buffer holding the result (or <0).
int radix_sort_3kernel(const KeySrc &src0, uint32_t n, int bits, Workspace &w, cudaStream_t stream)
{
    const uint32_t n_blocks = (n + kSortTile - 1) / kSortTile;
    const int passes = (bits + 7) / 8;
    int cur = -1;
    for (int p = 0; p < passes; ++p) {
        KeySrc src = src0;
        const uint32_t *vals_in = nullptr;
        if (p > 0) { src.keys = w.keys[cur]; src.from_ids = 0; vals_in = w.vals[cur]; }
        int dst = (cur + 1) & 1;
        if (cur < 0) dst = 0;
        k_radix_hist<<<n_blocks, kSortBlock, 0, stream>>>(src, n, p * 8, w.hist, n_blocks);
        k_radix_scan_rows<<<256, 1024, 0, stream>>>(w.hist, n_blocks, w.totals);
        k_radix_scan_totals<<<1, 256, 0, stream>>>(w.totals);
        k_radix_scatter<<<n_blocks, kSortBlock, 0, stream>>>(src, vals_in, p == 0 ? 1 : 0, n, p * 8, w.hist,
                                                             w.totals, n_blocks, w.keys[dst], w.vals[dst]);
        cur = dst;
    }
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return ts_cuda_fail(e, "radix sort launch");
    return cur;
}

int radix_onesweep_enabled()
{
    const char *e = getenv("TSRATE_ONESWEEP");   // read every call: tests switch it
    return e ? atoi(e) : 1;
}

// LSD radix sort of n (key, val) pairs on the low `bits` bits.  Pass 0 reads
// `src` (with identity values); results ping-pong between the two buffers.
// Returns the index of the buffer holding the result (or <0).
// staged: the pairs already stand in w.keys[0] / w.vals[0] and their number in *n_dev
// (n is then only an upper bound that sizes the grids): the owned slots of a sharded ladder.
// ext_vals (staged only): the pairs stand in src0.keys / ext_vals instead (memory the peers wrote).
int radix_sort(const KeySrc &src0, uint32_t n, int bits, Workspace &w, cudaStream_t stream, int sm_count,
               const uint32_t *n_dev = nullptr, const uint32_t *key_bound = nullptr, const uint32_t *ext_vals = nullptr)
{
    const int passes = (bits + 7) / 8;
    const bool staged = n_dev != nullptr;
    if (ext_vals && (!staged || passes < 1)) return TS_E_ARG;
    if (n >= (1u << 30) || passes > kMaxPasses || !radix_onesweep_enabled()) {
        if (key_bound) return TS_E_UNSUPPORTED;      // the caller sizes the sort on the host instead
        if (staged) {
            ts_set_error("the sharded ladder needs the single-sweep sort (fewer than 2^30 owned slot entries per window)");
            return TS_E_UNSUPPORTED;
        }
        return radix_sort_3kernel(src0, n, bits, w, stream);
    }
    const uint32_t n_tiles = (n + kSortTile - 1) / kSortTile;
    uint32_t *ghist = w.totals, *ticket = w.totals + kMaxPasses * 256;
    if (cudaMemsetAsync(ghist, 0, kMaxPasses * 256 * 4, stream) != cudaSuccess) return ts_cuda_fail(cudaGetLastError(), "memset");
    if (!staged && src0.from_ids && (src0.slots & 1u) == 0u && ((uintptr_t)src0.keys & 7u) == 0u)
        k_radix_hist_matches<<<sm_count * 8, kSortBlock, 0, stream>>>(src0, n / src0.slots, passes, ghist);
    else
        k_radix_hist_all<<<sm_count * 8, kSortBlock, 0, stream>>>(src0, n, n_dev, passes, ghist);
    k_radix_scan_all<<<1, 256, 0, stream>>>(ghist, passes);
    int cur = staged ? 0 : -1;
    for (int p = 0; p < passes; ++p) {
        KeySrc src = src0;
        const uint32_t *vals_in = nullptr;
        int dst = cur < 0 ? 0 : ((cur + 1) & 1);
        if (ext_vals && p == 0) { vals_in = ext_vals; dst = 0; }
        else if (cur >= 0) { src.keys = w.keys[cur]; src.from_ids = 0; vals_in = w.vals[cur]; }
        if (cudaMemsetAsync(w.hist, 0, (size_t)n_tiles * 256 * 4, stream) != cudaSuccess ||
            cudaMemsetAsync(ticket, 0, 4, stream) != cudaSuccess)
            return ts_cuda_fail(cudaGetLastError(), "memset");
        k_radix_onesweep<<<n_tiles, kSortBlock, 0, stream>>>(src, vals_in, vals_in ? 0 : 1, n, n_dev, p * 8, ghist + p * 256,
                                                             w.hist, ticket, w.keys[dst], w.vals[dst], key_bound);
        cur = dst;
    }
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return ts_cuda_fail(e, "radix sort launch");
    return cur;
}

uint32_t pred_range_entries();

// pred[v] for every slot of the sorted pairs in w.keys/vals[buf] (pred pre-filled by the caller).
// Destination ranges of at most 64 MiB of pred (16 Mi entries) per pass: half of the 126 MB L2.
// Every pass re-reads the sorted pairs, so this only pays for a handful of passes; a window
// too large for that (more than 128 Mi slot entries) takes one pass and the read-modify-writes.
void launch_pred(const Workspace &w, int buf, uint32_t np, const uint32_t *n_dev, const KeySrc &src, uint32_t seq_base,
                 const uint32_t *last, uint32_t *pred, cudaStream_t s)
{
    uint32_t range = pred_range_entries();
    if (((uint64_t)np + range - 1) / range > 8) range = 0xFFFFFFFFu;
    for (uint64_t lo = 0; lo < np; lo += range) {
        uint32_t hi = (uint32_t)std::min<uint64_t>(lo + range, np);
        k_pred<<<(np + 1023) / 1024, 256, 0, s>>>(w.keys[buf], w.vals[buf], np, n_dev, src.slots_magic, src.sentinel,
                                                  (uint32_t)lo, hi, seq_base, last, pred);
    }
}

int levels_blocks_per_sm_cap()
{
    static int cap = -1;
    if (cap < 0) {
        const char *e = getenv("TSRATE_LEVELS_BLOCKS_PER_SM");
        cap = e ? atoi(e) : 2;
        if (cap < 1) cap = 1;
    }
    return cap;
}

int schedule_frontier_enabled()
{
    const char *e = getenv("TSRATE_SCHEDULE_FRONTIER");   // read every call: tests switch it
    return e ? atoi(e) : 0;
}

}  // namespace

extern "C" int64_t ts_schedule_workspace_bytes(int64_t n_matches, int32_t slots)
{
    if (n_matches < 0 || slots < 2 || slots > TS_MAX_SLOTS) return TS_E_ARG;
    return (int64_t)carve(nullptr, n_matches, slots).bytes;
}

extern "C" int ts_schedule(ts_ctx *ctx, const uint32_t *ids, const uint32_t *meta, int64_t n_matches,
                           int32_t slots, int64_t n_players, const int8_t mode_col[16],
                           uint32_t *round, uint32_t *order, uint32_t *round_offsets, int64_t offsets_cap,
                           uint32_t *ids_sorted, uint32_t *meta_sorted, ts_sched_info *info,
                           void *workspace, int64_t workspace_bytes, uint32_t *n_rounds_host, void *stream)
{
    if (!ctx || !mode_col || !round_offsets || !info ||
        (n_matches > 0 && (!ids || !meta || !round || !order || !workspace))) {
        ts_set_error("ts_schedule: null argument");
        return TS_E_ARG;
    }
    if (slots < 2 || slots > TS_MAX_SLOTS || (slots & 1)) {
        ts_set_error("ts_schedule: slots must be even and in [2, %d] (got %d)", TS_MAX_SLOTS, slots);
        return TS_E_ARG;
    }
    if (n_matches < 0 || (uint64_t)n_matches * (uint64_t)slots >= 0xFFFF0000ull) {
        ts_set_error("ts_schedule: n_matches * slots must stay below 2^32 - 2^16: schedule the ladder in windows");
        return TS_E_ARG;
    }
    if (n_matches >= 0x7FFFFFFFll) {
        ts_set_error("ts_schedule: n_matches must stay below 2^31 - 1");
        return TS_E_ARG;
    }
    if (n_players < 1 || n_players >= 0xFFFFFFFFll) {
        ts_set_error("ts_schedule: n_players outside [1, 2^32 - 1)");
        return TS_E_ARG;
    }
    if (offsets_cap < 3) {
        ts_set_error("ts_schedule: round_offsets needs at least 3 entries");
        return TS_E_WORKSPACE;
    }
    TS_CUDA(cudaSetDevice(ctx->device));
    cudaStream_t s = (cudaStream_t)stream;
    if (n_matches == 0) {
        TS_CUDA(cudaMemsetAsync(info, 0, sizeof(ts_sched_info), s));
        TS_CUDA(cudaMemsetAsync(round_offsets, 0, 3 * sizeof(uint32_t), s));
        if (n_rounds_host) *n_rounds_host = 0;
        return 0;
    }
    Workspace w = carve((char *)workspace, n_matches, slots);
    if ((int64_t)w.bytes > workspace_bytes) {
        ts_set_error("ts_schedule: workspace too small (%lld < %lld bytes)", (long long)workspace_bytes, (long long)w.bytes);
        return TS_E_WORKSPACE;
    }
    const uint32_t n = (uint32_t)n_matches;
    const uint32_t np = n * (uint32_t)slots;
    ModeCols modes;
    for (int i = 0; i < 16; ++i) modes.col[i] = i < 6 ? mode_col[i] : (int8_t)-1;
    const uint32_t cap = (uint32_t)(offsets_cap > 0xFFFFFFFFll ? 0xFFFFFFFFll : offsets_cap);

    // 1. sort slots by player id
    KeySrc src;
    src.keys = ids;
    src.meta = meta;
    src.modes = modes;
    src.slots = (uint32_t)slots;
    src.sentinel = (uint32_t)n_players;
    src.from_ids = 1;
    src.slots_magic = (~0ull / (unsigned long long)slots) + 1ull;
    src.rated_mask = 0;
    for (int i = 0; i < 6; ++i) if (src.modes.col[i] >= 0) src.rated_mask |= 1u << i;
    int buf = radix_sort(src, np, bits_for((uint64_t)n_players), w, s, ctx->sm_count);
    if (buf < 0) return buf;

    // 2. + 3. rounds
    TS_CUDA(cudaMemsetAsync(round, 0, (size_t)n * 4, s));
    // barrier words and queue tail/snapshot; the error counter (word 2) is left alone
    if (int rc = ts_reset_barrier(ctx->d_ctrl, s)) return rc;
    TS_CUDA(cudaMemsetAsync(ctx->d_ctrl + CTRL_TAIL, 0, 2 * sizeof(unsigned long long), s));
    TS_CUDA(cudaMemsetAsync(ctx->d_ctrl + CTRL_LEVEL, 0, 3 * sizeof(unsigned long long), s));
    const bool by_levels = !schedule_frontier_enabled();
    if (by_levels) {
        // links to the previous match of every slot's player, then the rounds by dataflow along them
        uint32_t *pred = w.succ;
        TS_CUDA(cudaMemsetAsync(pred, 0xFF, (size_t)np * 4, s));
        launch_pred(w, buf, np, nullptr, src, 0u, nullptr, pred, s);
        TS_CUDA(cudaMemsetAsync(info, 0, sizeof(ts_sched_info), s));
        void *fn = slots == 6 ? (void *)k_levels<6> : slots == 10 ? (void *)k_levels<10> : (void *)k_levels<0>;
        int per_sm = 0;
        TS_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, fn, 256, 0));
        if (per_sm < 1) { ts_set_error("levels kernel does not fit"); return TS_E_CUDA; }
        // every resident lane polls: more than a few levels of matches in flight only floods L2
        per_sm = std::min(per_sm, levels_blocks_per_sm_cap());
        dim3 grid(ctx->sm_count * per_sm), block(256);
        uint32_t slots_u = (uint32_t)slots;
        const uint32_t *pred_c = pred;
        void *kargs[] = {&pred_c, &meta, &src.rated_mask, (void *)&n, &slots_u, &round, &info};
        // cooperative launch only for its guarantee that the whole grid is resident
        TS_CUDA(cudaLaunchCooperativeKernel(fn, grid, block, kargs, 0, s));
    } else {
        // level-synchronous frontier sweep (TSRATE_SCHEDULE_FRONTIER=1): successor links, one
        // persistent cooperative kernel, one grid-wide step per round
        k_link<<<(np + 255) / 256, 256, 0, s>>>(w.keys[buf], w.vals[buf], np, (uint32_t)slots, (uint32_t)n_players,
                                                w.succ);
        k_init_frontier<<<(n + 255) / 256, 256, 0, s>>>(meta, modes, w.succ, (uint32_t)slots, n, w.pending, w.queue,
                                                        ctx->d_ctrl);
        // close level 1: the sweep reads its first `end` from the snapshot word
        TS_CUDA(cudaMemcpyAsync(ctx->d_ctrl + CTRL_TAIL + 1, ctx->d_ctrl + CTRL_TAIL, sizeof(unsigned long long),
                                cudaMemcpyDeviceToDevice, s));
        int per_sm = 0;
        TS_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_frontier, 256, 0));
        if (per_sm < 1) { ts_set_error("frontier kernel does not fit"); return TS_E_CUDA; }
        per_sm = std::min(per_sm, frontier_blocks_per_sm_cap());
        dim3 grid(ctx->sm_count * per_sm), block(256);
        uint32_t slots_u = (uint32_t)slots;
        unsigned long long *ctrl = ctx->d_ctrl;
        void *kargs[] = {&w.queue, &w.succ, &w.pending, &round, &slots_u, (void *)&n, &round_offsets,
                         (void *)&cap, &info, &ctrl};
        TS_CUDA(cudaLaunchCooperativeKernel((void *)k_frontier, grid, block, kargs, 0, s));
        k_offsets<<<1, 1024, 0, s>>>(round_offsets, info, cap);
    }

    // 4. bucket by round, ascending match_seq inside a round
    KeySrc rsrc;
    rsrc.keys = round;
    rsrc.meta = nullptr;
    rsrc.modes = modes;
    rsrc.slots = 1;
    rsrc.sentinel = 0;
    rsrc.from_ids = 0;
    rsrc.slots_magic = 0ull;   // unused: keys are read as they are
    rsrc.rated_mask = 0;
    bool async = by_levels && !n_rounds_host && n < (1u << 30) && radix_onesweep_enabled();
    if (async) {
        // The number of rounds stays on the device: the sort is launched for as many passes as n
        // matches could need and the passes beyond the largest round skip themselves; bounds,
        // order and the overflow flag are written by k_round_bounds.  No host wait anywhere.
        buf = radix_sort(rsrc, n, bits_for((uint64_t)n), w, s, ctx->sm_count, nullptr, &info->n_rounds);
        if (buf < 0) return buf;
        k_round_bounds<<<n / 256 + 1, 256, 0, s>>>(w.keys[0], w.keys[1], w.vals[0], w.vals[1], n, info, cap, round_offsets,
                                                   order);
    } else {
        ts_sched_info h;
        TS_CUDA(cudaMemcpyAsync(&h, info, sizeof h, cudaMemcpyDeviceToHost, s));
        TS_CUDA(cudaStreamSynchronize(s));
        if (by_levels && (uint64_t)h.n_rounds + 2 > (uint64_t)cap) {
            h.overflow = 1;
            TS_CUDA(cudaMemcpyAsync(info, &h, sizeof h, cudaMemcpyHostToDevice, s));
            TS_CUDA(cudaStreamSynchronize(s));
        }
        if (h.overflow) {
            ts_set_error("ts_schedule: %u rounds need offsets_cap >= %u (got %lld)", h.n_rounds, h.n_rounds + 2,
                         (long long)offsets_cap);
            return TS_E_WORKSPACE;
        }
        buf = radix_sort(rsrc, n, bits_for((uint64_t)h.n_rounds), w, s, ctx->sm_count);
        if (buf < 0) return buf;
        if (by_levels)
            k_round_bounds<<<n / 256 + 1, 256, 0, s>>>(w.keys[buf], w.keys[buf], w.vals[buf], w.vals[buf], n, info, cap,
                                                       round_offsets, order);
        else
            TS_CUDA(cudaMemcpyAsync(order, w.vals[buf], (size_t)n * 4, cudaMemcpyDeviceToDevice, s));
        if (n_rounds_host) *n_rounds_host = h.n_rounds;
    }
    if (ids_sorted && meta_sorted) {
        size_t total = (size_t)n * slots;
        k_permute<<<(unsigned)((total + 255) / 256), 256, 0, s>>>(ids, meta, order, n, (uint32_t)slots, ids_sorted,
                                                                  meta_sorted);
    }
    TS_CUDA(cudaGetLastError());
    return 0;
}


extern "C" int64_t ts_link_workspace_bytes(int64_t n_matches, int32_t slots)
{
    if (n_matches < 0 || slots < 2 || slots > TS_MAX_SLOTS) return TS_E_ARG;
    size_t np = (size_t)n_matches * slots;
    size_t n_blocks = (np + kSortTile - 1) / kSortTile;
    return (int64_t)(4 * align_up(np * 4) + align_up((n_blocks * 256 + 256) * 4) + align_up(kTotalsWords * 4));
}

extern "C" int ts_link_pred(ts_ctx *ctx, const uint32_t *ids, const uint32_t *meta, int64_t n_matches, int32_t slots,
                            int64_t n_players, const int8_t mode_col[16], int64_t seq_base, uint32_t *last,
                            uint32_t *pred, void *workspace, int64_t workspace_bytes, void *stream)
{
    if (seq_base < 0 || seq_base + n_matches >= 0xFFFFFFFFll) {
        ts_set_error("ts_link_pred: stream-wide match numbers must stay below 2^32 - 1");
        return TS_E_ARG;
    }
    if (!ctx || !mode_col || (n_matches > 0 && (!ids || !meta || !pred || !workspace))) {
        ts_set_error("ts_link_pred: null argument");
        return TS_E_ARG;
    }
    if (slots < 2 || slots > TS_MAX_SLOTS || (slots & 1) || n_matches < 0 ||
        (uint64_t)n_matches * (uint64_t)slots >= 0xFFFF0000ull || n_players < 1 || n_players >= 0xFFFFFFFFll) {
        ts_set_error("ts_link_pred: bad sizes (n_matches * slots must stay below 2^32 - 2^16)");
        return TS_E_ARG;
    }
    if (n_matches == 0) return 0;
    if (workspace_bytes < ts_link_workspace_bytes(n_matches, slots)) {
        ts_set_error("ts_link_pred: workspace too small");
        return TS_E_WORKSPACE;
    }
    TS_CUDA(cudaSetDevice(ctx->device));
    cudaStream_t s = (cudaStream_t)stream;
    const uint32_t np = (uint32_t)n_matches * (uint32_t)slots;
    Workspace w = carve_link((char *)workspace, np);
    KeySrc src;
    src.keys = ids;
    src.meta = meta;
    for (int i = 0; i < 16; ++i) src.modes.col[i] = i < 6 ? mode_col[i] : (int8_t)-1;
    src.slots = (uint32_t)slots;
    src.sentinel = (uint32_t)n_players;
    src.from_ids = 1;
    src.slots_magic = (~0ull / (unsigned long long)slots) + 1ull;
    src.rated_mask = 0;
    for (int i = 0; i < 6; ++i) if (src.modes.col[i] >= 0) src.rated_mask |= 1u << i;
    int buf = radix_sort(src, np, bits_for((uint64_t)n_players), w, s, ctx->sm_count);
    if (buf < 0) return buf;
    // slots that take no part (empty, gated match) read TS_NONE: the output is fully defined
    TS_CUDA(cudaMemsetAsync(pred, 0xFF, (size_t)np * 4, s));
    launch_pred(w, buf, np, nullptr, src, (uint32_t)seq_base, last, pred, s);
    if (last)
        k_last_update<<<(np + 255) / 256, 256, 0, s>>>(w.keys[buf], w.vals[buf], np, nullptr, src.slots_magic,
                                                       src.sentinel, (uint32_t)seq_base, last);
    TS_CUDA(cudaGetLastError());
    return 0;
}

// ---------------------------------------------------------------------------
// Links for the slot-indexed form of the streaming rating kernel (csrc/slots.cu): succ[v] = slot
// index (m' * P + s') of the same player's next rated match, TS_NONE after the last; the head of
// every player's run gets the player's pair from the table as its row.  Same sorted runs as pred,
// looked at in the other direction; same L2-sized destination ranges for the scatter.
// ---------------------------------------------------------------------------
namespace {

__device__ __forceinline__ void st_row_gpu(unsigned long long *row, double mu, double sigma, uint32_t ver)
{
    const unsigned long long v = (unsigned long long)ver << 32;
    const unsigned long long m = (unsigned long long)__double_as_longlong(mu);
    const unsigned long long s = (unsigned long long)__double_as_longlong(sigma);
    asm volatile("st.relaxed.gpu.global.v4.u64 [%0], {%1, %2, %3, %4};" ::"l"(row), "l"((m & 0xFFFFFFFFull) | v),
                 "l"((m >> 32) | v), "l"((s & 0xFFFFFFFFull) | v), "l"((s >> 32) | v) : "memory");
}

__global__ void __launch_bounds__(256)
k_succ_slots(const uint32_t *keys, const uint32_t *vals, uint32_t n, unsigned long long slots_magic, uint32_t sentinel,
             uint32_t v_lo, uint32_t v_hi, int heads, const double *table, uint32_t epoch, uint32_t *succ,
             unsigned long long *rows, unsigned long long *ctrl)
{
    // four consecutive sorted entries per thread, plus the one before and the one after them
    const uint32_t i0 = (blockIdx.x * blockDim.x + threadIdx.x) * 4u;
    if (i0 >= n) return;
    uint32_t k[6], v[6];
#pragma unroll
    for (int j = 0; j < 6; ++j) {
        const int64_t i = (int64_t)i0 + j - 1;
        const bool in = i >= 0 && i < (int64_t)n;
        k[j] = in ? __ldcs(keys + i) : 0xFFFFFFFFu;
        v[j] = in ? __ldcs(vals + i) : 0u;
    }
#pragma unroll
    for (int j = 1; j <= 4; ++j) {
        if (i0 + j - 1 >= n || k[j] >= sentinel) continue;                  // empty slot / gated match / id outside the table
        uint32_t nxt = TS_NONE;
        if (i0 + j < n && k[j + 1] == k[j]) {
            nxt = v[j + 1];
            if ((uint32_t)__umul64hi((unsigned long long)nxt, slots_magic) == (uint32_t)__umul64hi((unsigned long long)v[j], slots_magic))
                atomicOr(&ctrl[CTRL_SHARD], (unsigned long long)TS_SHARD_DUPLICATE);   // the player twice in one match
        }
        if (v[j] >= v_lo && v[j] < v_hi) succ[v[j]] = nxt;
        if (heads && (i0 + j - 1 == 0 || k[j - 1] != k[j])) {               // the player's first match of the window
            const double2 pr = __ldg(reinterpret_cast<const double2 *>(table) + k[j]);
            st_row_gpu(rows + (size_t)v[j] * 4, pr.x, pr.y, epoch);
        }
    }
}

}  // namespace

extern "C" int ts_link_succ(ts_ctx *ctx, const uint32_t *ids, const uint32_t *meta, int64_t n_matches, int32_t slots,
                            const ts_table *table, uint32_t epoch, uint32_t *succ, void *rows, void *workspace,
                            int64_t workspace_bytes, void *stream)
{
    if (!ctx || !table || !table->data || (n_matches > 0 && (!ids || !meta || !succ || !rows || !workspace))) {
        ts_set_error("ts_link_succ: null argument");
        return TS_E_ARG;
    }
    if (table->row_pairs != 1 || table->shared_col >= 0) {
        ts_set_error("ts_link_succ: the table must be a single-pair ladder (row_pairs == 1, shared_col == -1)");
        return TS_E_UNSUPPORTED;
    }
    const int64_t n_players = table->n_players;
    if (slots < 2 || slots > TS_MAX_SLOTS || (slots & 1) || n_matches < 0 ||
        (uint64_t)n_matches * (uint64_t)slots >= 0xFFFF0000ull || n_players < 1 || n_players >= 0xFFFFFFFFll) {
        ts_set_error("ts_link_succ: bad sizes (n_matches * slots must stay below 2^32 - 2^16)");
        return TS_E_ARG;
    }
    if (epoch == 0u || epoch == TS_NONE || ((uintptr_t)rows & 31u) != 0u) {
        ts_set_error("ts_link_succ: epoch must be neither 0 nor 0xFFFFFFFF, rows 32-byte aligned");
        return TS_E_ARG;
    }
    if (n_matches == 0) return 0;
    if (workspace_bytes < ts_link_workspace_bytes(n_matches, slots)) {
        ts_set_error("ts_link_succ: workspace too small");
        return TS_E_WORKSPACE;
    }
    TS_CUDA(cudaSetDevice(ctx->device));
    cudaStream_t s = (cudaStream_t)stream;
    const uint32_t np = (uint32_t)n_matches * (uint32_t)slots;
    Workspace w = carve_link((char *)workspace, np);
    KeySrc src;
    src.keys = ids;
    src.meta = meta;
    for (int i = 0; i < 16; ++i) src.modes.col[i] = i < 6 ? table->mode_col[i] : (int8_t)-1;
    src.slots = (uint32_t)slots;
    src.sentinel = (uint32_t)n_players;
    src.from_ids = 1;
    src.slots_magic = (~0ull / (unsigned long long)slots) + 1ull;
    src.rated_mask = 0;
    for (int i = 0; i < 6; ++i) if (src.modes.col[i] >= 0) src.rated_mask |= 1u << i;
    int buf = radix_sort(src, np, bits_for((uint64_t)n_players), w, s, ctx->sm_count);
    if (buf < 0) return buf;
    uint32_t range = pred_range_entries();
    if (((uint64_t)np + range - 1) / range > 8) range = 0xFFFFFFFFu;
    for (uint64_t lo = 0; lo < np; lo += range) {
        const uint32_t hi = (uint32_t)std::min<uint64_t>(lo + range, np);
        k_succ_slots<<<(np + 1023) / 1024, 256, 0, s>>>(w.keys[buf], w.vals[buf], np, src.slots_magic, src.sentinel, (uint32_t)lo,
                                                       hi, lo == 0 ? 1 : 0, table->data, epoch, succ,
                                                       (unsigned long long *)rows, ctx->d_ctrl);
    }
    TS_CUDA(cudaGetLastError());
    return 0;
}

// ---------------------------------------------------------------------------
// Links of ONE ladder spread over the G GPUs of a node (tsrate.h: ts_shard_view).
// Every rank links the players it owns (p % G == rank):
//   1. k_owned_stage / k_tile_scan / k_owned_gather read the whole id stream once -- the own
//      slice and, over NVLink, the peers' -- and keep, in stream order, (local player index,
//      slot index of the window) of every live slot of an owned player;
//   2. the pairs are radix-sorted by local player index (same single-sweep passes as the
//      one-GPU sort, their number known only on the device);
//   3. k_succ_route walks the sorted runs: a slot's successor match becomes a message to the
//      rank that rates the slot's match, appended to that rank's inbox in large coalesced
//      pieces (the receiver scatters them into its succ slice, k_succ_apply), and the head of
//      every run sends the player's pair from the owner's table shard to the mailbox of the
//      rank that rates the player's first match, tagged with that match's number.
// ---------------------------------------------------------------------------
namespace {

struct ShardSrc {
    const uint32_t *ids[TS_SHARD_MAX];
    const uint32_t *meta[TS_SHARD_MAX];
    uint32_t g_mask, g_shift;       // G - 1, log2 G
    uint32_t owner;                 // this rank
    uint32_t slots;
    unsigned long long slots_magic;
    uint32_t rated_mask;
    uint32_t n_players;
};

// slot idx = m * slots + s of the window -> local index of its player on this rank, or TS_NONE
__device__ __forceinline__ uint32_t shard_key(const ShardSrc &src, uint32_t idx)
{
    const uint32_t m = (uint32_t)__umul64hi((unsigned long long)idx, src.slots_magic);
    const uint32_t sl = idx - m * src.slots;
    const uint32_t blk = m >> 5, r = blk & src.g_mask;
    const uint32_t lm = ((blk >> src.g_shift) << 5) | (m & 31u);
    const uint32_t k = src.ids[r][(size_t)lm * src.slots + sl];
    if (k == TS_EMPTY_SLOT || k >= src.n_players || (k & src.g_mask) != src.owner) return TS_NONE;
    const uint32_t mt = src.meta[r][lm];
    const bool gated = ((src.rated_mask >> ((mt >> TS_META_MODE_SHIFT) & 15u)) & 1u) == 0u ||
                       (mt & (TS_META_AFK | TS_META_INVALID)) != 0u;
    return gated ? TS_NONE : (k >> src.g_shift);
}

// Compaction of the owned slots in three independent steps -- no tile waits for another:
//   k_owned_stage   every tile of 4096 slots compacts its owned (player, slot) pairs to the head of
//                   its own 4096-entry piece of a staging array and stores their number;
//   k_tile_scan     exclusive scan of the tile counts (one block);
//   k_owned_gather  every tile copies its piece to its place in the sort's input.
// (A single pass with decoupled look-back between the tiles was built first: half of its time went
// to waiting for the tiles before -- ncu: 2.7 ms for 300M slots, issue slots 23 % used.)
__global__ void __launch_bounds__(kSortBlock)
k_owned_stage(ShardSrc src, uint32_t np, unsigned long long *staging, uint32_t *counts)
{
    __shared__ uint32_t warp_tot[kWarps];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const uint32_t tile = blockIdx.x;
    const uint32_t wbase = tile * kSortTile + warp * (kSortItems * 32);
    uint32_t key[kSortItems], off[kSortItems];
#pragma unroll
    for (int k = 0; k < kSortItems; ++k) {
        const uint32_t idx = wbase + k * 32 + lane;
        key[k] = idx < np ? shard_key(src, idx) : TS_NONE;
    }
    uint32_t running = 0;
#pragma unroll
    for (int k = 0; k < kSortItems; ++k) {
        const uint32_t bal = __ballot_sync(0xFFFFFFFFu, key[k] != TS_NONE);
        off[k] = running + __popc(bal & ((1u << lane) - 1u));
        running += __popc(bal);
    }
    if (lane == 0) warp_tot[warp] = running;
    __syncthreads();
    uint32_t before = 0, total = 0;
#pragma unroll
    for (int w = 0; w < kWarps; ++w) { before += (w < warp) ? warp_tot[w] : 0u; total += warp_tot[w]; }
    if (threadIdx.x == 0) counts[tile] = total;
    unsigned long long *piece = staging + (size_t)tile * kSortTile;
#pragma unroll
    for (int k = 0; k < kSortItems; ++k)
        if (key[k] != TS_NONE) piece[before + off[k]] = ((unsigned long long)(wbase + k * 32 + lane) << 32) | key[k];
}

// in place: counts[t] := number of owned slots in the tiles before t; counts[n_tiles] := their total
__global__ void __launch_bounds__(1024)
k_tile_scan(uint32_t *counts, uint32_t n_tiles, uint32_t cap, uint32_t *n_out, unsigned long long *ctrl)
{
    __shared__ uint32_t warp_sums[32];
    __shared__ uint32_t carry;
    if (threadIdx.x == 0) carry = 0;
    __syncthreads();
    for (uint32_t base = 0; base < n_tiles; base += 1024) {
        const uint32_t i = base + threadIdx.x;
        const uint32_t x = i < n_tiles ? counts[i] : 0u;
        uint32_t incl = x;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const uint32_t y = __shfl_up_sync(0xFFFFFFFFu, incl, o);
            if ((threadIdx.x & 31) >= o) incl += y;
        }
        if ((threadIdx.x & 31) == 31) warp_sums[threadIdx.x >> 5] = incl;
        __syncthreads();
        if (threadIdx.x < 32) {
            const uint32_t w = warp_sums[threadIdx.x];
            uint32_t wi = w;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const uint32_t y = __shfl_up_sync(0xFFFFFFFFu, wi, o);
                if (threadIdx.x >= o) wi += y;
            }
            warp_sums[threadIdx.x] = wi - w;
        }
        __syncthreads();
        const uint32_t excl = carry + warp_sums[threadIdx.x >> 5] + incl - x;
        if (i < n_tiles) counts[i] = excl;
        __syncthreads();
        if (threadIdx.x == 1023) carry = excl + x;
        __syncthreads();
    }
    if (threadIdx.x == 0) {
        counts[n_tiles] = carry;
        *n_out = carry < cap ? carry : cap;
        if (carry > cap) atomicOr(&ctrl[CTRL_SHARD], (unsigned long long)TS_SHARD_OVERFLOW);
    }
}

__global__ void __launch_bounds__(kSortBlock)
k_owned_gather(const unsigned long long *staging, const uint32_t *offsets, uint32_t cap, uint32_t *keys_out, uint32_t *vals_out)
{
    const uint32_t tile = blockIdx.x;
    const uint32_t base = __ldg(offsets + tile), cnt = __ldg(offsets + tile + 1) - base;
    const unsigned long long *piece = staging + (size_t)tile * kSortTile;
    for (uint32_t i = threadIdx.x; i < cnt; i += kSortBlock) {
        const unsigned long long e = __ldcs(piece + i);
        if (base + i < cap) { keys_out[base + i] = (uint32_t)e; vals_out[base + i] = (uint32_t)(e >> 32); }
    }
}

struct ShardDst {
    unsigned long long *inbox[TS_SHARD_MAX];     // rank g's inbox: [n_ranges][G senders][region_cap] messages
    uint32_t *inbox_count[TS_SHARD_MAX];         // rank g's message counts, [n_ranges][G senders]
    unsigned long long *rows[TS_SHARD_MAX];
    const double *table_own;
    uint32_t *out_count;                         // this rank's running message count per bucket (destination, range)
    uint32_t region_cap;                         // messages one sender may leave in one (destination, range) region
    uint32_t n_ranges, range_shift;              // a destination's slots in ranges of 2^range_shift: one L2-sized scatter each
    uint32_t g_mask, g_shift, owner;
    uint32_t slots;
    unsigned long long slots_magic;
    uint32_t seq_base;
};

__device__ __forceinline__ void st_row_sys(unsigned long long *row, double mu, double sigma, uint32_t ver)
{
    const unsigned long long v = (unsigned long long)ver << 32;
    const unsigned long long m = (unsigned long long)__double_as_longlong(mu);
    const unsigned long long s = (unsigned long long)__double_as_longlong(sigma);
    asm volatile("st.relaxed.sys.global.v4.u64 [%0], {%1, %2, %3, %4};" ::"l"(row), "l"((m & 0xFFFFFFFFull) | v),
                 "l"((m >> 32) | v), "l"((s & 0xFFFFFFFFull) | v), "l"((s >> 32) | v) : "memory");
}

// Successor links, routed: a scattered 4-byte store into another GPU's memory is a partial-sector
// write over NVLink (measured on 2 B200s: 150M of them take 185 ms, 0.8 G/s; scripts/nvlink_probe.py),
// so the links do not go to their slots directly.  A block of kRouteTile sorted entries sorts its
// (local slot index, successor) messages by bucket = (destination rank, range of the destination's
// slots) in shared memory and appends each bucket's run to this sender's region of that rank's inbox
// with consecutive threads on consecutive addresses (full sectors, large NVLink packets); the
// destination scatters its inbox locally, one L2-sized range at a time (k_succ_apply), so every
// sector of succ goes to HBM once, complete.  The head of every player's run also sends the
// player's pair from the owner's table shard to the mailbox of the rank that rates the player's
// first match (a full 32-byte row).
constexpr int kRouteItems = 8, kRouteTile = 256 * kRouteItems, kRouteBuckets = 256;

__global__ void __launch_bounds__(256)
k_succ_route(const uint32_t *keys, const uint32_t *vals, const uint32_t *n_dev, ShardDst dst, unsigned long long *ctrl)
{
    __shared__ uint32_t s_cnt[kRouteBuckets], s_start[kRouteBuckets + 1], s_base[kRouteBuckets];
    __shared__ unsigned long long s_msg[kRouteTile];
    const uint32_t n = __ldg(n_dev);
    const uint32_t tile_base = blockIdx.x * kRouteTile;
    if (tile_base >= n) return;
    const uint32_t n_buckets = (dst.g_mask + 1u) * dst.n_ranges;
    s_cnt[threadIdx.x] = 0;
    __syncthreads();
    uint32_t off[kRouteItems], to[kRouteItems];
    unsigned long long msg[kRouteItems];
#pragma unroll
    for (int j = 0; j < kRouteItems; ++j) {
        const uint32_t i = tile_base + j * 256 + threadIdx.x;
        to[j] = TS_NONE;
        if (i >= n) continue;
        const uint32_t k = __ldcs(keys + i), v = __ldcs(vals + i);
        const uint32_t m = (uint32_t)__umul64hi((unsigned long long)v, dst.slots_magic);
        const uint32_t sl = v - m * dst.slots;
        uint32_t nxt = TS_NONE;
        if (i + 1 < n && __ldcs(keys + i + 1) == k) {
            const uint32_t m2 = (uint32_t)__umul64hi((unsigned long long)__ldcs(vals + i + 1), dst.slots_magic);
            if (m2 == m) atomicOr(&ctrl[CTRL_SHARD], (unsigned long long)TS_SHARD_DUPLICATE);
            nxt = dst.seq_base + m2;
        }
        const uint32_t blk = m >> 5, r = blk & dst.g_mask;
        const uint32_t lm = ((blk >> dst.g_shift) << 5) | (m & 31u);
        const uint32_t lv = lm * dst.slots + sl;
        to[j] = r * dst.n_ranges + min(lv >> dst.range_shift, dst.n_ranges - 1u);
        msg[j] = ((unsigned long long)nxt << 32) | (unsigned long long)lv;
        off[j] = atomicAdd(&s_cnt[to[j]], 1u);
        if (i == 0 || __ldcs(keys + i - 1) != k) {            // the player's first match of the window
            const double2 pr = __ldg(reinterpret_cast<const double2 *>(dst.table_own) + k);
            const uint32_t pid = (k << dst.g_shift) | dst.owner;
            st_row_sys(dst.rows[r] + (size_t)pid * 4, pr.x, pr.y, dst.seq_base + m);
        }
    }
    __syncthreads();
    if (threadIdx.x < n_buckets) {
        const uint32_t c = s_cnt[threadIdx.x];
        s_base[threadIdx.x] = c ? atomicAdd(&dst.out_count[threadIdx.x], c) : 0u;   // this block's place in every region
    }
    if (threadIdx.x < 32) {                                    // exclusive scan of the bucket counts, one warp
        uint32_t run = 0;
        for (uint32_t b0 = 0; b0 < n_buckets; b0 += 32) {
            const uint32_t b = b0 + threadIdx.x;
            const uint32_t c = b < n_buckets ? s_cnt[b] : 0u;
            uint32_t incl = c;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const uint32_t y = __shfl_up_sync(0xFFFFFFFFu, incl, o);
                if (threadIdx.x >= (unsigned)o) incl += y;
            }
            if (b < n_buckets) s_start[b] = run + incl - c;
            run += __shfl_sync(0xFFFFFFFFu, incl, 31);
        }
        if (threadIdx.x == 0) s_start[n_buckets] = run;
    }
    __syncthreads();
#pragma unroll
    for (int j = 0; j < kRouteItems; ++j)
        if (to[j] != TS_NONE) s_msg[s_start[to[j]] + off[j]] = msg[j];
    __syncthreads();
    const uint32_t total = min((uint32_t)kRouteTile, n - tile_base);
    for (uint32_t pos = threadIdx.x; pos < total; pos += 256) {
        uint32_t lo = 0, hi = n_buckets;                       // the bucket whose run holds pos
        while (hi - lo > 1) {
            const uint32_t mid = (lo + hi) >> 1;
            if (s_start[mid] <= pos) lo = mid; else hi = mid;
        }
        const uint32_t idx = s_base[lo] + (pos - s_start[lo]);
        const uint32_t g = lo / dst.n_ranges, range = lo - g * dst.n_ranges;
        if (idx < dst.region_cap)
            dst.inbox[g][((size_t)range * (dst.g_mask + 1u) + dst.owner) * dst.region_cap + idx] = s_msg[pos];
        else atomicOr(&ctrl[CTRL_SHARD], (unsigned long long)TS_SHARD_OVERFLOW);
    }
}

// after all messages are out: tell every rank how many this sender left in each of its regions
__global__ void k_succ_counts(ShardDst dst)
{
    const uint32_t b = threadIdx.x;
    if (b >= (dst.g_mask + 1u) * dst.n_ranges) return;
    const uint32_t g = b / dst.n_ranges, range = b - g * dst.n_ranges;
    dst.inbox_count[g][range * (dst.g_mask + 1u) + dst.owner] = min(dst.out_count[b], dst.region_cap);
}

// the receiving side, local: succ[local slot] = successor for every message of every sender in one range
__global__ void __launch_bounds__(256)
k_succ_apply(const unsigned long long *regions, const uint32_t *counts, uint32_t region_cap, uint32_t *succ)
{
    const uint32_t sender = blockIdx.y;
    const uint32_t n = __ldg(counts + sender);
    const unsigned long long *src = regions + (size_t)sender * region_cap;
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        const unsigned long long msg = __ldcs(src + i);
        succ[(uint32_t)msg] = (uint32_t)(msg >> 32);
    }
}

// How a rank's succ slice of `entries` slots is cut into ranges: 2^shift slots each, 16 Mi (64 MiB,
// half an L2) unless that makes more buckets than the routing kernel sorts.
void shard_ranges(uint64_t entries, int n_shards, uint32_t *n_ranges, uint32_t *shift)
{
    uint32_t sh = 24;
    while (((entries + (1ull << sh) - 1) >> sh) * (uint64_t)n_shards > (uint64_t)kRouteBuckets) ++sh;
    *shift = sh;
    *n_ranges = (uint32_t)std::max<uint64_t>(1, (entries + (1ull << sh) - 1) >> sh);
}

// ---------------------------------------------------------------------------
// The owned (player, slot) pairs WITHOUT every rank reading the whole id stream: every rank looks
// at its own slice only and sends each pair to the player's owner, to the exact place it has in
// stream order there.  Per block of 32 matches and owner the number of pairs is counted
// (k_block_owned<false>), the counts of all ranks are exchanged (host: one small all-gather), an
// exclusive scan over the blocks in GLOBAL order (k_chunk_sums / k_chunk_scan / k_block_offsets)
// gives every (block, owner) its offset in the owner's input, and k_block_owned<true> stores the
// pairs there -- consecutive pairs of a block to consecutive addresses of the owner's memory, a
// few hundred bytes at a time over NVLink.  Link work per rank ~ 1/G of the stream.
// ---------------------------------------------------------------------------
struct ExchSrc {
    const uint32_t *ids, *meta;     // own slice
    uint32_t n_local;               // matches in it
    uint32_t g_mask, g_shift, me;
    uint32_t slots, rated_mask, n_players;
    unsigned long long slots_magic; // floor(2^64 / slots) + 1
    uint32_t nb_cap;                // blocks per rank the count / offset arrays are laid out for
};
struct ExchDst {
    uint32_t *in_keys[TS_SHARD_MAX], *in_vals[TS_SHARD_MAX];
    uint32_t in_cap;
};
struct ExchPeers {
    uint4 *counts[TS_SHARD_MAX];    // every rank's link_counts at sender = me
};

// owner (and local index) of the player in entry e = match * slots + slot of block lb of the own slice; 0xFF: takes no part
__device__ __forceinline__ uint32_t exch_owner(const ExchSrc &src, uint32_t lb, uint32_t e, uint32_t &key, uint32_t &val)
{
    const uint32_t mi = (uint32_t)__umul64hi((unsigned long long)e, src.slots_magic), sl = e - mi * src.slots;
    const uint32_t lm = lb * 32u + mi;
    if (lm >= src.n_local) return 0xFFu;
    const uint32_t k = __ldg(src.ids + (size_t)lm * src.slots + sl);
    if (k == TS_EMPTY_SLOT || k >= src.n_players) return 0xFFu;
    const uint32_t mt = __ldg(src.meta + lm);
    const bool gated = ((src.rated_mask >> ((mt >> TS_META_MODE_SHIFT) & 15u)) & 1u) == 0u ||
                       (mt & (TS_META_AFK | TS_META_INVALID)) != 0u;
    if (gated) return 0xFFu;
    key = k >> src.g_shift;
    // slot index of the window: the stream-wide number of local match lm of this rank
    val = (((((lm >> 5) << src.g_shift) | src.me) << 5) | (lm & 31u)) * src.slots + sl;
    return k & src.g_mask;
}

// One WARP = one block of 32 matches of the own slice (32 * slots entries, 32 at a time); lane o keeps
// owner o's count.  No block-wide step.  SEND: a second sweep puts the block's pairs, grouped by owner,
// into the warp's piece of shared memory, so that the pairs for one owner leave as one run of
// consecutive 4-byte stores (about 32 * slots / G of them) to that owner's memory.
constexpr int kExchWarps = 4;
template <bool SEND>
__global__ void __launch_bounds__(kExchWarps * 32)
k_block_owned(ExchSrc src, uint32_t nb_local, uint16_t *cnt_local, const uint32_t *off_mine, ExchDst dst)
{
    extern __shared__ uint4 exch_smem[];                    // SEND: [kExchWarps][32 * slots] of (key, val, place, owner)
    const uint32_t warp = threadIdx.x >> 5, lane = threadIdx.x & 31, lt = (1u << lane) - 1u;
    const uint32_t lb = blockIdx.x * kExchWarps + warp;
    if (lb >= nb_local) return;
    const uint32_t G = src.g_mask + 1u;
    // (loading all 32 * slots entries before the counting, unrolled over TS_MAX_SLOTS, measured slower:
    // 80 registers, send 0.21 -> 0.28 ms for 2.5M matches)
    uint32_t tot = 0;                                       // lane o: pairs of this block for owner o
    for (uint32_t c = 0; c < src.slots; ++c) {
        uint32_t key, val;
        const uint32_t owner = exch_owner(src, lb, c * 32u + lane, key, val);
        for (uint32_t o = 0; o < G; ++o) {
            const uint32_t bal = __ballot_sync(0xFFFFFFFFu, owner == o);
            if (lane == o) tot += __popc(bal);
        }
    }
    if (!SEND) {
        if (lane < G) cnt_local[(size_t)lane * src.nb_cap + lb] = (uint16_t)tot;
        return;
    }
    uint32_t incl = tot;                                    // start of every owner's run in the warp's piece
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const uint32_t y = __shfl_up_sync(0xFFFFFFFFu, incl, o);
        if (lane >= (uint32_t)o) incl += y;
    }
    const uint32_t seg = incl - tot, total = __shfl_sync(0xFFFFFFFFu, incl, 31);
    const uint32_t my_off = lane < G ? __ldg(off_mine + (size_t)lane * src.nb_cap + lb) : 0u;
    uint4 *piece = exch_smem + (size_t)warp * 32u * src.slots;
    uint32_t run = 0;                                       // lane o: owner o's pairs in the entries before
    for (uint32_t c = 0; c < src.slots; ++c) {
        uint32_t key = 0, val = 0, idx = 0;
        const uint32_t owner = exch_owner(src, lb, c * 32u + lane, key, val);
        for (uint32_t o = 0; o < G; ++o) {
            const uint32_t bal = __ballot_sync(0xFFFFFFFFu, owner == o);
            const uint32_t before = __shfl_sync(0xFFFFFFFFu, run, o);
            if (owner == o) idx = before + __popc(bal & lt);
            if (lane == o) run += __popc(bal);
        }
        const uint32_t sg = __shfl_sync(0xFFFFFFFFu, seg, owner & 31u), off = __shfl_sync(0xFFFFFFFFu, my_off, owner & 31u);
        if (owner != 0xFFu) piece[sg + idx] = make_uint4(key, val, off + idx, owner);
    }
    __syncwarp();
    for (uint32_t j = lane; j < total; j += 32u) {
        const uint4 e = piece[j];
        if (e.z < dst.in_cap) { dst.in_keys[e.w][e.z] = e.x; dst.in_vals[e.w][e.z] = e.y; }
    }
}

// own counts [G owners][nb_cap] -> the same place (sender = me) in every peer's link_counts
__global__ void __launch_bounds__(256) k_push_counts(const uint4 *own, ExchPeers peers, uint32_t me, uint32_t G, size_t n16)
{
    for (size_t i = (size_t)blockIdx.x * 256 + threadIdx.x; i < n16; i += (size_t)gridDim.x * 256) {
        const uint4 v = own[i];
        for (uint32_t g = 0; g < G; ++g)
            if (g != me) peers.counts[g][i] = v;
    }
}

// cnt_all[sender][owner][nb_cap] (16-bit); global block b = lb * G + sender.  A chunk = 4096 global blocks.
constexpr uint32_t kExchChunk = 4096;

__device__ __forceinline__ uint32_t exch_count(const uint16_t *cnt_all, uint32_t b, uint32_t owner, uint32_t g_mask,
                                               uint32_t g_shift, uint32_t nb_cap, uint32_t n_blocks)
{
    if (b >= n_blocks) return 0u;
    const uint32_t sender = b & g_mask, lb = b >> g_shift;
    return cnt_all[((size_t)sender * (g_mask + 1u) + owner) * nb_cap + lb];
}

__global__ void __launch_bounds__(256)
k_chunk_sums(const uint16_t *cnt_all, uint32_t g_mask, uint32_t g_shift, uint32_t nb_cap, uint32_t n_blocks, uint32_t n_chunks,
             uint32_t *partial)
{
    __shared__ uint32_t red[8];
    const uint32_t owner = blockIdx.y, chunk = blockIdx.x;
    uint32_t sum = 0;
    for (uint32_t i = threadIdx.x; i < kExchChunk; i += 256)
        sum += exch_count(cnt_all, chunk * kExchChunk + i, owner, g_mask, g_shift, nb_cap, n_blocks);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) sum += __shfl_xor_sync(0xFFFFFFFFu, sum, o);
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = sum;
    __syncthreads();
    if (threadIdx.x == 0) {
        uint32_t tot = 0;
        for (int w = 0; w < 8; ++w) tot += red[w];
        partial[(size_t)owner * n_chunks + chunk] = tot;
    }
}

// one block per owner: exclusive scan of its chunk sums in place; the owner's total to totals[owner]
// (rank `me` also stores its total, clamped to in_cap, as the number of pairs it receives)
__global__ void __launch_bounds__(1024) k_chunk_scan(uint32_t *partial, uint32_t n_chunks, uint32_t *totals, uint32_t me,
                                                     uint32_t in_cap, uint32_t *in_count, unsigned long long *ctrl)
{
    __shared__ uint32_t warp_sums[32];
    __shared__ uint32_t carry;
    uint32_t *row = partial + (size_t)blockIdx.x * n_chunks;
    if (threadIdx.x == 0) carry = 0;
    __syncthreads();
    for (uint32_t base = 0; base < n_chunks; base += 1024) {
        const uint32_t i = base + threadIdx.x;
        const uint32_t x = i < n_chunks ? row[i] : 0u;
        uint32_t incl = x;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const uint32_t y = __shfl_up_sync(0xFFFFFFFFu, incl, o);
            if ((threadIdx.x & 31) >= o) incl += y;
        }
        if ((threadIdx.x & 31) == 31) warp_sums[threadIdx.x >> 5] = incl;
        __syncthreads();
        if (threadIdx.x < 32) {
            const uint32_t w = warp_sums[threadIdx.x];
            uint32_t wi = w;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const uint32_t y = __shfl_up_sync(0xFFFFFFFFu, wi, o);
                if (threadIdx.x >= o) wi += y;
            }
            warp_sums[threadIdx.x] = wi - w;
        }
        __syncthreads();
        const uint32_t excl = carry + warp_sums[threadIdx.x >> 5] + incl - x;
        if (i < n_chunks) row[i] = excl;
        __syncthreads();
        if (threadIdx.x == 1023) carry = excl + x;
        __syncthreads();
    }
    if (threadIdx.x == 0) {
        totals[blockIdx.x] = carry;
        if (blockIdx.x == me) {
            *in_count = min(carry, in_cap);
            if (carry > in_cap) atomicOr(&ctrl[CTRL_SHARD], (unsigned long long)TS_SHARD_OVERFLOW);
        }
    }
}

// offsets of this rank's own blocks: off_mine[owner][lb] = pairs for `owner` in all global blocks before (lb, me)
__global__ void __launch_bounds__(256)
k_block_offsets(const uint16_t *cnt_all, const uint32_t *partial, uint32_t g_mask, uint32_t g_shift, uint32_t me,
                uint32_t nb_cap, uint32_t n_blocks, uint32_t n_chunks, uint32_t *off_mine)
{
    __shared__ uint32_t warp_sums[8];
    const uint32_t owner = blockIdx.y, chunk = blockIdx.x;
    constexpr uint32_t per = kExchChunk / 256;             // consecutive global blocks per thread
    const uint32_t b0 = chunk * kExchChunk + threadIdx.x * per;
    uint32_t c[per], sum = 0;
#pragma unroll
    for (uint32_t j = 0; j < per; ++j) { c[j] = exch_count(cnt_all, b0 + j, owner, g_mask, g_shift, nb_cap, n_blocks); sum += c[j]; }
    uint32_t incl = sum;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const uint32_t y = __shfl_up_sync(0xFFFFFFFFu, incl, o);
        if ((threadIdx.x & 31) >= (unsigned)o) incl += y;
    }
    if ((threadIdx.x & 31) == 31) warp_sums[threadIdx.x >> 5] = incl;
    __syncthreads();
    uint32_t run = partial[(size_t)owner * n_chunks + chunk] + incl - sum;
    for (uint32_t w = 0; w < (threadIdx.x >> 5); ++w) run += warp_sums[w];
#pragma unroll
    for (uint32_t j = 0; j < per; ++j) {
        const uint32_t b = b0 + j;
        if (b < n_blocks && (b & g_mask) == me) off_mine[(size_t)owner * nb_cap + (b >> g_shift)] = run;
        run += c[j];
    }
}

// the number of pairs ts_shard_send left for this rank, bounded by what the sort's buffers hold
__global__ void k_take_count(const uint32_t *in_count, uint32_t cap, uint32_t *n_out, unsigned long long *ctrl)
{
    const uint32_t n = *in_count;
    *n_out = min(n, cap);
    if (n > cap) atomicOr(&ctrl[CTRL_SHARD], (unsigned long long)TS_SHARD_OVERFLOW);
}

int check_view(const ts_shard_view *v, const char *who)
{
    if (!v || v->n_shards < 1 || v->n_shards > TS_SHARD_MAX || (v->n_shards & (v->n_shards - 1)) || v->rank < 0 ||
        v->rank >= v->n_shards) {
        ts_set_error("%s: n_shards must be a power of two <= %d and 0 <= rank < n_shards", who, TS_SHARD_MAX);
        return TS_E_ARG;
    }
    return 0;
}

}  // namespace

// the local half of the link exchange (called by ts_shard_rate before the rating kernel): rank `me`
// scatters the messages its inbox holds into its succ slice
int ts_shard_apply_links(const ts_shard_view *view, int me, int64_t n_matches, int32_t slots, cudaStream_t s)
{
    const int G = view->n_shards;
    uint32_t n_ranges = 1, shift = 24;
    shard_ranges((uint64_t)ts_shard_local_matches(n_matches, 0, G) * (uint64_t)slots, G, &n_ranges, &shift);
    const uint32_t region_cap = (uint32_t)(view->inbox_cap / n_ranges);
    const unsigned blocks = (unsigned)std::min<uint32_t>((region_cap + 1023u) / 1024u, 2048u);
    for (uint32_t r = 0; r < n_ranges; ++r)
        k_succ_apply<<<dim3(blocks, (unsigned)G), 256, 0, s>>>(view->inbox[me] + (size_t)r * G * region_cap,
                                                            view->inbox_count[me] + (size_t)r * G, region_cap, view->succ[me]);
    TS_CUDA(cudaGetLastError());
    return 0;
}

// Messages one sender may leave in one receiver's inbox for a window of n_matches: per range of the
// receiver's slots its share of the range (a G-th, with room for uneven ownership), times the ranges.
extern "C" int64_t ts_shard_inbox_cap(int64_t n_matches, int32_t slots, int32_t n_shards)
{
    if (n_matches < 0 || slots < 2 || n_shards < 1 || n_shards > TS_SHARD_MAX) return TS_E_ARG;
    const uint64_t entries = (uint64_t)ts_shard_local_matches(n_matches, 0, n_shards) * (uint64_t)slots;
    uint32_t n_ranges = 1, shift = 24;
    shard_ranges(entries, n_shards, &n_ranges, &shift);
    const uint64_t per_range = std::min<uint64_t>(entries, 1ull << shift);
    const uint64_t region = n_shards == 1 ? per_range : std::min<uint64_t>(per_range, per_range * 13 / (10 * (uint64_t)n_shards) + 4096);
    return (int64_t)(std::max<uint64_t>(region, 1) * n_ranges + kRouteBuckets);
}

extern "C" int64_t ts_shard_local_matches(int64_t n_matches, int32_t rank, int32_t n_shards)
{
    if (n_matches < 0 || n_shards < 1 || rank < 0 || rank >= n_shards) return TS_E_ARG;
    const int64_t blocks = (n_matches + TS_SHARD_BLOCK - 1) / TS_SHARD_BLOCK;
    if (blocks <= rank) return 0;
    const int64_t mine = (blocks - rank + n_shards - 1) / n_shards;
    const bool last_is_mine = (blocks - 1) % n_shards == rank;
    return (mine - 1) * TS_SHARD_BLOCK + (last_is_mine ? n_matches - (blocks - 1) * TS_SHARD_BLOCK : TS_SHARD_BLOCK);
}

extern "C" int64_t ts_shard_link_workspace_bytes(int64_t n_slot_entries, int64_t owned_cap)
{
    if (owned_cap < 0 || owned_cap >= 0xFFFF0000ll || n_slot_entries < 0 || n_slot_entries >= 0xFFFF0000ll) return TS_E_ARG;
    const int64_t tiles = (n_slot_entries + kSortTile - 1) / kSortTile;
    return ts_link_workspace_bytes(owned_cap / 2 + 1, 2) + (int64_t)align_up((size_t)tiles * kSortTile * 8) +
           (int64_t)align_up((size_t)(tiles + 1) * 4);
}

namespace {

// TSRATE_SHARD_TIMING=1: the time of one entry point's launches (synchronises; diagnosis only)
struct StageTimer {
    cudaEvent_t e0 = nullptr, e1 = nullptr;
    cudaStream_t s;
    const char *what;
    int rank, G;
    StageTimer(const char *w, int r, int g, cudaStream_t st) : s(st), what(w), rank(r), G(g)
    {
        const char *e = getenv("TSRATE_SHARD_TIMING");
        if (e && atoi(e)) { cudaEventCreate(&e0); cudaEventCreate(&e1); cudaEventRecord(e0, s); }
    }
    ~StageTimer()
    {
        if (!e0) return;
        cudaEventRecord(e1, s);
        cudaEventSynchronize(e1);
        float ms = 0;
        cudaEventElapsedTime(&ms, e0, e1);
        fprintf(stderr, "%s rank %d/%d: %.3f ms\n", what, rank, G, ms);
        cudaEventDestroy(e0);
        cudaEventDestroy(e1);
    }
};

uint32_t exch_blocks(int64_t n_matches) { return (uint32_t)((n_matches + TS_SHARD_BLOCK - 1) / TS_SHARD_BLOCK); }

int check_exchange(const ts_shard_view *view, int64_t n_matches, int32_t slots, int64_t n_players, int mode, const char *who)
{
    if (int rc = check_view(view, who)) return rc;
    const int G = view->n_shards;
    if (n_matches < 0 || slots < 2 || slots > TS_MAX_SLOTS || (slots & 1) || n_players < 1 || n_players >= 0xFFFFFFFFll ||
        (uint64_t)n_matches * (uint64_t)slots >= 0xFFFF0000ull || mode < 0 || mode > 5) {
        ts_set_error("%s: bad sizes (n_matches * slots must stay below 2^32 - 2^16, mode 0..5)", who);
        return TS_E_ARG;
    }
    if (view->link_blocks_cap < 8 || (view->link_blocks_cap & 7) || view->in_cap < 1 || view->in_cap >= 0xFFFF0000ll ||
        (int64_t)(exch_blocks(n_matches) + G - 1) / G > view->link_blocks_cap) {
        ts_set_error("%s: link_blocks_cap (a multiple of 8) or in_cap too small for the window", who);
        return TS_E_ARG;
    }
    for (int g = 0; g < G; ++g)
        if (!view->link_counts[g] || !view->in_keys[g] || !view->in_vals[g] || !view->ids[view->rank] || !view->meta[view->rank] ||
            !view->in_count[view->rank]) {
            ts_set_error("%s: shard %d's link exchange arrays not mapped", who, g);
            return TS_E_ARG;
        }
    return 0;
}

ExchSrc exchange_src(const ts_shard_view *view, int64_t n_matches, int32_t slots, int64_t n_players, int mode)
{
    const int G = view->n_shards, rank = view->rank;
    int shift = 0;
    while ((1 << shift) < G) ++shift;
    ExchSrc src;
    src.ids = view->ids[rank];
    src.meta = view->meta[rank];
    src.n_local = (uint32_t)ts_shard_local_matches(n_matches, rank, G);
    src.g_mask = (uint32_t)G - 1u;
    src.g_shift = (uint32_t)shift;
    src.me = (uint32_t)rank;
    src.slots = (uint32_t)slots;
    src.slots_magic = (~0ull / (unsigned long long)slots) + 1ull;
    src.rated_mask = 1u << mode;
    src.n_players = (uint32_t)n_players;
    src.nb_cap = (uint32_t)view->link_blocks_cap;
    return src;
}

}  // namespace

extern "C" int64_t ts_shard_exchange_workspace_bytes(int64_t n_matches, int32_t n_shards)
{
    if (n_matches < 0 || n_shards < 1 || n_shards > TS_SHARD_MAX) return TS_E_ARG;
    const int64_t n_blocks = exch_blocks(n_matches), n_chunks = (n_blocks + kExchChunk - 1) / kExchChunk;
    const int64_t nb_cap = ((n_blocks + n_shards - 1) / n_shards + 7) & ~7ll;
    return (int64_t)align_up((size_t)n_shards * (size_t)(n_chunks + 1) * 4) + 256 + (int64_t)align_up((size_t)n_shards * (size_t)nb_cap * 4);
}

extern "C" int ts_shard_count(ts_ctx *ctx, const ts_shard_view *view, int64_t n_matches, int32_t slots, int64_t n_players,
                              int8_t mode, void *stream)
{
    if (!ctx) return TS_E_ARG;
    if (int rc = check_exchange(view, n_matches, slots, n_players, mode, "ts_shard_count")) return rc;
    if (n_matches == 0) return 0;
    TS_CUDA(cudaSetDevice(ctx->device));
    cudaStream_t s = (cudaStream_t)stream;
    const int G = view->n_shards, rank = view->rank;
    const ExchSrc src = exchange_src(view, n_matches, slots, n_players, mode);
    StageTimer timer("ts_shard_count", rank, G, s);
    const size_t per_sender = (size_t)G * src.nb_cap;               // 16-bit counts of one sender
    uint16_t *own = view->link_counts[rank] + (size_t)rank * per_sender;
    const uint32_t nb_local = (src.n_local + 31u) / 32u;
    if (nb_local) k_block_owned<false><<<(nb_local + kExchWarps - 1) / kExchWarps, kExchWarps * 32, 0, s>>>(src, nb_local, own, nullptr, ExchDst());
    if (G > 1) {
        ExchPeers peers;
        for (int g = 0; g < TS_SHARD_MAX; ++g)
            peers.counts[g] = g < G ? reinterpret_cast<uint4 *>(view->link_counts[g] + (size_t)rank * per_sender) : nullptr;
        const size_t n16 = per_sender * 2 / 16;
        k_push_counts<<<(unsigned)std::min<size_t>((n16 + 255) / 256, (size_t)ctx->sm_count * 4), 256, 0, s>>>(
            reinterpret_cast<const uint4 *>(own), peers, (uint32_t)rank, (uint32_t)G, n16);
    }
    TS_CUDA(cudaGetLastError());
    return 0;
}

extern "C" int ts_shard_send(ts_ctx *ctx, const ts_shard_view *view, int64_t n_matches, int32_t slots, int64_t n_players,
                             int8_t mode, void *workspace, int64_t workspace_bytes, void *stream)
{
    if (!ctx) return TS_E_ARG;
    if (int rc = check_exchange(view, n_matches, slots, n_players, mode, "ts_shard_send")) return rc;
    if (n_matches == 0) return 0;
    const int G = view->n_shards, rank = view->rank;
    const uint32_t n_blocks = exch_blocks(n_matches), n_chunks = (n_blocks + kExchChunk - 1) / kExchChunk;
    const ExchSrc src = exchange_src(view, n_matches, slots, n_players, mode);
    const size_t partial_bytes = align_up((size_t)G * (n_chunks + 1) * 4), off_bytes = align_up((size_t)G * src.nb_cap * 4);
    if (!workspace || ((uintptr_t)workspace & 15u) || (int64_t)(partial_bytes + 256 + off_bytes) > workspace_bytes) {
        ts_set_error("ts_shard_send: workspace too small (ts_shard_exchange_workspace_bytes) or not 16-byte aligned");
        return TS_E_WORKSPACE;
    }
    TS_CUDA(cudaSetDevice(ctx->device));
    cudaStream_t s = (cudaStream_t)stream;
    StageTimer timer("ts_shard_send", rank, G, s);
    uint32_t *partial = (uint32_t *)workspace;
    uint32_t *totals = (uint32_t *)((char *)workspace + partial_bytes);
    uint32_t *off_mine = (uint32_t *)((char *)workspace + partial_bytes + 256);
    const uint16_t *cnt_all = view->link_counts[rank];
    k_chunk_sums<<<dim3(n_chunks, (unsigned)G), 256, 0, s>>>(cnt_all, src.g_mask, src.g_shift, src.nb_cap, n_blocks, n_chunks, partial);
    k_chunk_scan<<<(unsigned)G, 1024, 0, s>>>(partial, n_chunks, totals, (uint32_t)rank, (uint32_t)view->in_cap, view->in_count[rank],
                                              ctx->d_ctrl);
    k_block_offsets<<<dim3(n_chunks, (unsigned)G), 256, 0, s>>>(cnt_all, partial, src.g_mask, src.g_shift, (uint32_t)rank, src.nb_cap,
                                                                n_blocks, n_chunks, off_mine);
    ExchDst dst;
    for (int g = 0; g < TS_SHARD_MAX; ++g) { dst.in_keys[g] = g < G ? view->in_keys[g] : nullptr; dst.in_vals[g] = g < G ? view->in_vals[g] : nullptr; }
    dst.in_cap = (uint32_t)view->in_cap;
    const uint32_t nb_local = (src.n_local + 31u) / 32u;
    if (nb_local)
        k_block_owned<true><<<(nb_local + kExchWarps - 1) / kExchWarps, kExchWarps * 32, (size_t)kExchWarps * 32 * slots * sizeof(uint4), s>>>(
            src, nb_local, nullptr, off_mine, dst);
    TS_CUDA(cudaGetLastError());
    return 0;
}

extern "C" int ts_shard_link(ts_ctx *ctx, const ts_shard_view *view, int64_t n_matches, int32_t slots, int64_t n_players,
                             int8_t mode, int64_t seq_base, void *workspace, int64_t workspace_bytes, void *stream)
{
    if (int rc = check_view(view, "ts_shard_link")) return rc;
    const int G = view->n_shards, rank = view->rank;
    if (!ctx || n_matches < 0 || slots < 2 || slots > TS_MAX_SLOTS || (slots & 1) || n_players < 1 ||
        n_players >= 0xFFFFFFFFll || (uint64_t)n_matches * (uint64_t)slots >= 0xFFFF0000ull || mode < 0 || mode > 5) {
        ts_set_error("ts_shard_link: bad sizes (n_matches * slots must stay below 2^32 - 2^16, mode 0..5)");
        return TS_E_ARG;
    }
    if (seq_base < 0 || seq_base % (TS_SHARD_BLOCK * G) != 0 || seq_base + n_matches >= 0xFFFFFFFFll) {
        ts_set_error("ts_shard_link: seq_base must be a multiple of 32 * n_shards and match numbers stay below 2^32 - 1");
        return TS_E_ARG;
    }
    if (n_matches == 0) return 0;
    for (int g = 0; g < G; ++g)
        if (!view->ids[g] || !view->meta[g] || !view->inbox[g] || !view->inbox_count[g] || !view->rows[g] || !view->table[g]) {
            ts_set_error("ts_shard_link: shard %d not mapped", g);
            return TS_E_ARG;
        }
    if (view->inbox_cap < kRouteBuckets || view->inbox_cap > 0x7FFFFFFFll) {
        ts_set_error("ts_shard_link: inbox_cap out of range");
        return TS_E_ARG;
    }
    // workspace: the sort's buffers for `cap` owned pairs (ts_link_workspace_bytes), then the staging
    // array of the compaction (8 bytes per slot of the window) and its tile counts
    // (link exchange, view->in_keys set: the pairs were sent here by ts_shard_send -- no staging)
    const bool exchanged = view->in_keys[rank] != nullptr;
    if (exchanged && (!view->in_vals[rank] || !view->in_count[rank] || view->in_cap < 1)) {
        ts_set_error("ts_shard_link: in_keys without in_vals / in_count / in_cap");
        return TS_E_ARG;
    }
    const uint32_t np = (uint32_t)n_matches * (uint32_t)slots;
    const int64_t compact_tiles = exchanged ? 0 : ((int64_t)np + kSortTile - 1) / kSortTile;
    const int64_t stage_bytes = (int64_t)align_up((size_t)compact_tiles * kSortTile * 8) + (int64_t)align_up((size_t)(compact_tiles + 1) * 4);
    int64_t cap = exchanged ? std::min<int64_t>(np, view->in_cap) : np;
    while (cap > 0 && ts_link_workspace_bytes(cap / 2 + 1, 2) + stage_bytes > workspace_bytes) cap -= (cap + 15) / 16;
    if (cap < 1 || !workspace) {
        ts_set_error("ts_shard_link: workspace too small");
        return TS_E_WORKSPACE;
    }
    TS_CUDA(cudaSetDevice(ctx->device));
    cudaStream_t s = (cudaStream_t)stream;
    // TSRATE_SHARD_TIMING=1: synchronise after every stage and print its time (diagnosis only)
    const bool timing = getenv("TSRATE_SHARD_TIMING") && atoi(getenv("TSRATE_SHARD_TIMING"));
    cudaEvent_t tev[4];
    if (timing) { for (auto &e : tev) cudaEventCreate(&e); cudaEventRecord(tev[0], s); }
    Workspace w = carve_link((char *)workspace, (uint32_t)cap);
    unsigned long long *staging = (unsigned long long *)((char *)workspace + w.bytes);
    uint32_t *tile_counts = (uint32_t *)((char *)staging + align_up((size_t)compact_tiles * kSortTile * 8));
    uint32_t *ticket = w.totals + kMaxPasses * 256, *n_dev = ticket + 8;
    int shift = 0;
    while ((1 << shift) < G) ++shift;
    ShardSrc src;
    for (int g = 0; g < TS_SHARD_MAX; ++g) { src.ids[g] = g < G ? view->ids[g] : nullptr; src.meta[g] = g < G ? view->meta[g] : nullptr; }
    src.g_mask = (uint32_t)G - 1u;
    src.g_shift = (uint32_t)shift;
    src.owner = (uint32_t)rank;
    src.slots = (uint32_t)slots;
    src.slots_magic = (~0ull / (unsigned long long)slots) + 1ull;
    src.rated_mask = 1u << mode;
    src.n_players = (uint32_t)n_players;
    TS_CUDA(cudaMemsetAsync(w.totals, 0, kTotalsWords * 4, s));
    if (exchanged) {
        k_take_count<<<1, 1, 0, s>>>(view->in_count[rank], (uint32_t)cap, n_dev, ctx->d_ctrl);
    } else {
        k_owned_stage<<<(unsigned)compact_tiles, kSortBlock, 0, s>>>(src, np, staging, tile_counts);
        k_tile_scan<<<1, 1024, 0, s>>>(tile_counts, (uint32_t)compact_tiles, (uint32_t)cap, n_dev, ctx->d_ctrl);
        k_owned_gather<<<(unsigned)compact_tiles, kSortBlock, 0, s>>>(staging, tile_counts, (uint32_t)cap, w.keys[0], w.vals[0]);
    }
    if (timing) cudaEventRecord(tev[1], s);
    const int64_t local_players = (n_players - rank + G - 1) / G;
    KeySrc ks;
    ks.keys = exchanged ? view->in_keys[rank] : w.keys[0];
    ks.meta = nullptr;
    for (int i = 0; i < 16; ++i) ks.modes.col[i] = (int8_t)-1;
    ks.slots = 1;
    ks.sentinel = 0xFFFFFFFFu;
    ks.from_ids = 0;
    ks.slots_magic = 0ull;
    ks.rated_mask = 0;
    int buf = radix_sort(ks, (uint32_t)cap, std::max(1, bits_for((uint64_t)(local_players > 0 ? local_players - 1 : 0))), w, s,
                         ctx->sm_count, n_dev, nullptr, exchanged ? view->in_vals[rank] : nullptr);
    if (buf < 0) return buf;
    if (timing) cudaEventRecord(tev[2], s);
    ShardDst dst;
    for (int g = 0; g < TS_SHARD_MAX; ++g) {
        dst.inbox[g] = g < G ? view->inbox[g] : nullptr;
        dst.inbox_count[g] = g < G ? view->inbox_count[g] : nullptr;
        dst.rows[g] = g < G ? view->rows[g] : nullptr;
    }
    dst.table_own = view->table[rank];
    // ranges of the largest slice (the slices differ by one block of 32 matches at most); same on every rank
    shard_ranges((uint64_t)ts_shard_local_matches(n_matches, 0, G) * (uint64_t)slots, G, &dst.n_ranges, &dst.range_shift);
    dst.region_cap = (uint32_t)(view->inbox_cap / dst.n_ranges);
    dst.out_count = w.hist;                      // the sort's tile status rows are free again
    TS_CUDA(cudaMemsetAsync(dst.out_count, 0, kRouteBuckets * 4, s));
    dst.g_mask = src.g_mask;
    dst.g_shift = src.g_shift;
    dst.owner = src.owner;
    dst.slots = src.slots;
    dst.slots_magic = src.slots_magic;
    dst.seq_base = (uint32_t)seq_base;
    k_succ_route<<<(unsigned)((cap + kRouteTile - 1) / kRouteTile), 256, 0, s>>>(w.keys[buf], w.vals[buf], n_dev, dst, ctx->d_ctrl);
    k_succ_counts<<<1, kRouteBuckets, 0, s>>>(dst);
    TS_CUDA(cudaGetLastError());
    if (timing) {
        cudaEventRecord(tev[3], s);
        cudaEventSynchronize(tev[3]);
        float ms[3];
        for (int i = 0; i < 3; ++i) cudaEventElapsedTime(&ms[i], tev[i], tev[i + 1]);
        uint32_t owned = 0;
        cudaMemcpy(&owned, n_dev, 4, cudaMemcpyDeviceToHost);
        fprintf(stderr, "ts_shard_link rank %d/%d: compact %.3f ms, sort %.3f ms, route %.3f ms (%u owned of %u slots, cap %lld)\n",
                rank, G, ms[0], ms[1], ms[2], owned, np, (long long)cap);
        for (auto &e : tev) cudaEventDestroy(e);
    }
    return 0;
}
