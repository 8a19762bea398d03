// Synthetic ladder generator (SURVEY.md 8d "Generator"): counter-based
// splitmix64 hashing, integer-only decisions, so the CUDA output is
// bit-identical to the host twin in oracle/ts_oracle.c.  Used by bench.py to
// build BASELINE.json's configs directly in HBM.  No reference counterpart.
#include "common.cuh"

namespace {

__host__ __device__ __forceinline__ uint64_t mix64(uint64_t z)
{
    z += 0x9E3779B97F4A7C15ull;
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
}
__device__ __forceinline__ uint64_t rnd3(uint64_t seed, uint64_t a, uint64_t b, uint64_t c)
{
    return mix64(mix64(mix64(seed ^ a) + b * 0xD1342543DE82EF95ull) + c);
}
__device__ __forceinline__ int32_t four_u16(uint64_t h)
{
    return (int32_t)((h & 0xFFFF) + ((h >> 16) & 0xFFFF) + ((h >> 32) & 0xFFFF) + (h >> 48)) - 2 * 65535;
}
__device__ __forceinline__ int32_t latent_skill(uint64_t seed, uint64_t pid)
{
    return four_u16(mix64(mix64(seed ^ 0x5EEDFACE5EEDFACEull) + pid));
}
__device__ __forceinline__ uint32_t sample_player(const ts_gen &g, uint64_t x)
{
    if (!g.zipf) return (uint32_t)__umul64hi(x, g.n_players);
    // inverse CDF of (rank + q)^-0.8 with explicitly rounded steps (no FMA contraction)
    double u = __dmul_rn((double)(x >> 11), 0x1.0p-53);
    double y = __dadd_rn(__dmul_rn(u, g.zipf_b), g.zipf_a);
    double y2 = __dmul_rn(y, y);
    double y4 = __dmul_rn(y2, y2);
    double y5 = __dmul_rn(y4, y);
    double r = __dadd_rn(y5, -g.zipf_q);
    if (!(r >= 0.0)) r = 0.0;
    uint64_t k = (uint64_t)r;
    if (k >= g.n_players) k = g.n_players - 1;
    return (uint32_t)k;
}

__global__ void __launch_bounds__(256) k_gen_matches(ts_gen g, int64_t first, int64_t count, uint32_t *ids, uint32_t *meta)
{
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= count) return;
    const uint64_t m = (uint64_t)(first + i);
    const int P = g.slots, half = P / 2;
    uint32_t pid[TS_MAX_SLOTS];
    int64_t strength = 0;
    for (int s = 0; s < P; ++s) {
        uint64_t draw = 0;
        for (;;) {
            uint32_t p = sample_player(g, rnd3(g.seed, m, (uint64_t)s, draw));
            bool dup = false;
            for (int k = 0; k < s; ++k) dup |= (pid[k] == p);
            if (!dup) { pid[s] = p; break; }
            ++draw;
        }
        int32_t sk = latent_skill(g.seed, pid[s]);
        strength += (s < half) ? sk : -sk;
        ids[i * P + s] = pid[s];
    }
    uint64_t x = rnd3(g.seed, m, 0xFFFFull, 0);
    long long thr = (1ll << 31) + strength * 11585ll;
    const long long lo = 85899346ll, hi = 4209067950ll;
    thr = thr < lo ? lo : (thr > hi ? hi : thr);
    uint32_t bits = ((long long)(x >> 32) < thr) ? TS_META_WIN_A : TS_META_WIN_B;
    uint64_t y = rnd3(g.seed, m, 0xFFFEull, 0);
    if ((uint32_t)y < g.draw_thr) bits = 0;
    if ((uint32_t)(y >> 32) < g.afk_thr) bits |= TS_META_AFK;
    meta[i] = bits | ((uint32_t)g.mode << TS_META_MODE_SHIFT);
}

__global__ void __launch_bounds__(256) k_gen_table(uint64_t seed, int64_t first, int64_t count, double *table)
{
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= count) return;
    uint64_t h = mix64(mix64(seed ^ 0x7AB1E7AB1E7AB1E5ull) + (uint64_t)(first + i));
    double z = __dmul_rn((double)four_u16(h), 1.0 / 37837.0);
    double mu = __dadd_rn(__dmul_rn(400.0, z), 1500.0);
    reinterpret_cast<double2 *>(table)[i] = make_double2(mu, 500.0 * (2.0 / 3.0));
}

}  // namespace

extern "C" int ts_gen_matches(const ts_gen *gen, int64_t first, int64_t count, uint32_t *ids, uint32_t *meta,
                              void *stream)
{
    if (!gen || !ids || !meta || count < 0 || gen->slots < 2 || gen->slots > TS_MAX_SLOTS || (gen->slots & 1) ||
        gen->n_players < (uint64_t)gen->slots) {
        ts_set_error("ts_gen_matches: bad argument");
        return TS_E_ARG;
    }
    if (count == 0) return 0;
    k_gen_matches<<<(unsigned)((count + 255) / 256), 256, 0, (cudaStream_t)stream>>>(*gen, first, count, ids, meta);
    TS_CUDA(cudaGetLastError());
    return 0;
}

extern "C" int ts_gen_table(uint64_t seed, int64_t first, int64_t count, double *table, void *stream)
{
    if (!table || count < 0) {
        ts_set_error("ts_gen_table: bad argument");
        return TS_E_ARG;
    }
    if (count == 0) return 0;
    k_gen_table<<<(unsigned)((count + 255) / 256), 256, 0, (cudaStream_t)stream>>>(seed, first, count, table);
    TS_CUDA(cudaGetLastError());
    return 0;
}

// ---------------------------------------------------------------------------
// Seed policy on the device: get_trueskill_seed (rater.py:42-62) for a whole
// column of players.  rank_points_* are doubles with NaN = None; 0 counts as
// absent like in the reference (rater.py:45,47).  vst[tier + 1] is the
// reference's vst_points table (rater.py:13-27), passed in by the host so that
// there is one source for it.  A tier outside -1..29 with no rank points raises
// KeyError in the reference: here the seed stays NaN and bad[i] = 1.
// ---------------------------------------------------------------------------
namespace {

struct VstTable {
    double pts[31];
};

__global__ void __launch_bounds__(256)
k_seed(const double *ranked, const double *blitz, const int8_t *tier, int64_t n, double unknown_sigma, VstTable vst,
       double *seed, uint8_t *bad)
{
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    double best = nan("");
    const double r = ranked ? ranked[i] : nan(""), b = blitz ? blitz[i] : nan("");
    if (!isnan(r) && r != 0.0) best = r;
    if (!isnan(b) && b != 0.0 && (isnan(best) || b > best)) best = b;
    double mu = nan(""), sigma = nan("");
    uint8_t flag = 0;
    if (!isnan(best)) {
        sigma = unknown_sigma * (2.0 / 3.0);
        mu = best + sigma;
    } else {
        const int t = tier ? (int)tier[i] : 127;
        if (t >= -1 && t <= 29) {
            sigma = unknown_sigma;
            mu = vst.pts[t + 1] + sigma;
        } else {
            flag = 1;
        }
    }
    reinterpret_cast<double2 *>(seed)[i] = make_double2(mu, sigma);
    if (bad) bad[i] = flag;
}

}  // namespace

extern "C" int ts_seed_table(const double *rank_points_ranked, const double *rank_points_blitz, const int8_t *skill_tier,
                             int64_t n_players, double unknown_player_sigma, const double vst_points[31], double *seed,
                             uint8_t *bad, void *stream)
{
    if (!seed || !vst_points || n_players < 0) {
        ts_set_error("ts_seed_table: bad argument");
        return TS_E_ARG;
    }
    if (n_players == 0) return 0;
    VstTable v;
    for (int i = 0; i < 31; ++i) v.pts[i] = vst_points[i];
    k_seed<<<(unsigned)((n_players + 255) / 256), 256, 0, (cudaStream_t)stream>>>(
        rank_points_ranked, rank_points_blitz, skill_tier, n_players, unknown_player_sigma, v, seed, bad);
    TS_CUDA(cudaGetLastError());
    return 0;
}

// Ingest helper: the meta word uses its low 8 bits only, so a host may ship it as bytes
// (a quarter of the PCIe volume of that column) and widen it here.
__global__ void __launch_bounds__(256) k_widen_meta(const uint8_t *meta8, int64_t n, uint32_t *meta)
{
    // four bytes per thread: one 4-byte load, one 16-byte store
    const int64_t i = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) * 4;
    if (i + 4 <= n && ((reinterpret_cast<uintptr_t>(meta8 + i) & 3u) == 0u) && ((reinterpret_cast<uintptr_t>(meta + i) & 15u) == 0u)) {
        const uint32_t w = __ldg(reinterpret_cast<const uint32_t *>(meta8 + i));
        *reinterpret_cast<uint4 *>(meta + i) = make_uint4(w & 255u, (w >> 8) & 255u, (w >> 16) & 255u, w >> 24);
        return;
    }
    for (int64_t j = i; j < n && j < i + 4; ++j) meta[j] = meta8[j];
}

extern "C" int ts_widen_meta(const uint8_t *meta8, int64_t n_matches, uint32_t *meta, void *stream)
{
    if (n_matches < 0 || (n_matches > 0 && (!meta8 || !meta))) {
        ts_set_error("ts_widen_meta: bad arguments");
        return TS_E_ARG;
    }
    if (n_matches == 0) return 0;
    const int64_t threads = (n_matches + 3) / 4;
    k_widen_meta<<<(unsigned)((threads + 255) / 256), 256, 0, (cudaStream_t)stream>>>(meta8, n_matches, meta);
    TS_CUDA(cudaGetLastError());
    return 0;
}

// Ingest helper: player ids shipped at the width they need.  A host column of n ids packed as a
// little-endian bit stream of `bits` bits each (bits = ceil(log2(n_players + 1)); the all-ones
// value stands for TS_EMPTY_SLOT) is 20/32 of the 32-bit column for a million players -- and the
// PCIe copy of that column is what bounds the end-to-end call.  Four ids per thread: the 4 * bits
// bits they occupy start on a nibble boundary at worst, read as unaligned bytes.
__global__ void __launch_bounds__(256) k_unpack_ids(const uint8_t *packed, int bits, int64_t count, uint32_t *ids)
{
    const int64_t i0 = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) * 4;
    if (i0 >= count) return;
    const uint32_t mask = bits >= 32 ? 0xFFFFFFFFu : ((1u << bits) - 1u);
    uint32_t out[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
        const int64_t i = i0 + j;
        out[j] = TS_EMPTY_SLOT;
        if (i >= count) continue;
        const int64_t bit = i * bits;
        const uint8_t *src = packed + (bit >> 3);
        unsigned long long w = 0;                       // up to 5 bytes hold 7 + 32 bits
        const int n_bytes = (int)(((bit & 7) + bits + 7) >> 3);
        for (int b = 0; b < n_bytes; ++b) w |= (unsigned long long)__ldg(src + b) << (8 * b);
        const uint32_t v = (uint32_t)(w >> (bit & 7)) & mask;
        out[j] = v == mask ? TS_EMPTY_SLOT : v;
    }
    if (i0 + 4 <= count && ((reinterpret_cast<uintptr_t>(ids + i0) & 15u) == 0u)) {
        *reinterpret_cast<uint4 *>(ids + i0) = make_uint4(out[0], out[1], out[2], out[3]);
        return;
    }
    for (int j = 0; j < 4 && i0 + j < count; ++j) ids[i0 + j] = out[j];
}

extern "C" int ts_unpack_ids(const uint8_t *packed, int32_t bits, int64_t count, uint32_t *ids, void *stream)
{
    if (bits < 1 || bits > 32 || count < 0 || (count > 0 && (!packed || !ids))) {
        ts_set_error("ts_unpack_ids: bad arguments (bits must be 1..32)");
        return TS_E_ARG;
    }
    if (count == 0) return 0;
    const int64_t threads = (count + 3) / 4;
    k_unpack_ids<<<(unsigned)((threads + 255) / 256), 256, 0, (cudaStream_t)stream>>>(packed, bits, count, ids);
    TS_CUDA(cudaGetLastError());
    return 0;
}
