// Host side of the columnar ingest: the participant rows a worker's query yields (worker.py:176-191,
// one row per participant, ORDER BY created_at) -> the columns the kernels read.  Plain C++ on the
// host (no device code): two passes over the rows with two small hash maps, where the numpy form of
// the same thing (analyzer_b200/rater.py: ingest_rows) spends a millisecond in library calls for a
// 500-match batch.  Mirrors rater.py:42-62 (seed policy), 70-106 (mode, roster-count and AFK gates).
#include <math.h>
#include <stdint.h>
#include <string.h>

#include <vector>

#include "common.cuh"

namespace {

// int64 key -> dense index in order of first appearance (open addressing, power-of-two table that doubles
// at half load: its size follows the number of distinct keys, not the number of rows -- a table laid out for
// as many keys as rows cost more to clear than the whole batch to ingest; key and index share a 16-byte
// entry, and the caller prefetches the entry of a row a few rows ahead)
struct FirstSeen {
    struct Entry { int64_t key; int32_t val; int32_t pad; };
    std::vector<Entry> tab;
    uint64_t mask;
    int32_t count = 0;
    explicit FirstSeen(int64_t rows)          // small batches never grow; large ones start at 64k entries
    {
        uint64_t cap = 1024;
        while (cap < (uint64_t)rows * 2 + 2 && cap < (1u << 16)) cap <<= 1;
        tab.assign(cap, Entry{0, -1, 0});
        mask = cap - 1;
    }
    static uint64_t hash(int64_t k)
    {
        uint64_t h = (uint64_t)k * 0x9E3779B97F4A7C15ull;
        return h ^ (h >> 29);
    }
    void prefetch(int64_t k) const { __builtin_prefetch(&tab[hash(k) & mask]); }
    void grow()
    {
        std::vector<Entry> old(tab.size() * 2, Entry{0, -1, 0});
        old.swap(tab);
        mask = tab.size() - 1;
        for (const Entry &e : old) {
            if (e.val < 0) continue;
            uint64_t i = hash(e.key) & mask;
            while (tab[i].val >= 0) i = (i + 1) & mask;
            tab[i] = e;
        }
    }
    int32_t get(int64_t k, bool *fresh)
    {
        for (uint64_t i = hash(k) & mask;; i = (i + 1) & mask) {
            Entry &e = tab[i];
            if (e.key == k && e.val >= 0) { *fresh = false; return e.val; }
            if (e.val < 0) {
                e.key = k; e.val = count; *fresh = true;
                const int32_t idx = count++;
                if ((uint64_t)count * 2 > mask) grow();
                return idx;
            }
        }
    }
};

}  // namespace

extern "C" int ts_ingest_rows(const ts_row_columns *rows, double unknown_sigma, const double vst_points[31],
                              int64_t *match_keys, int64_t *player_keys, int32_t *row_match, int32_t *row_slot,
                              int32_t *row_player, uint32_t *ids, uint32_t *meta, double *table, double *seed,
                              int64_t *n_matches_out, int64_t *n_players_out, int32_t *slots_out)
{
    if (!rows || rows->n_rows < 0 || !vst_points || !match_keys || !player_keys || !row_match || !row_slot || !row_player ||
        !ids || !meta || !table || !seed || !n_matches_out || !n_players_out || !slots_out) {
        ts_set_error("ts_ingest_rows: null argument");
        return TS_E_ARG;
    }
    const int64_t n = rows->n_rows;
    if (n >= 0x7FFFFFFFll) { ts_set_error("ts_ingest_rows: fewer than 2^31 rows per batch"); return TS_E_ARG; }
    if (n > 0 && (!rows->match || !rows->player || !rows->game_mode || !rows->n_rosters || !rows->roster || !rows->winner ||
                  !rows->went_afk || !rows->skill_tier || !rows->rank_points_ranked || !rows->rank_points_blitz ||
                  !rows->trueskill_mu || !rows->trueskill_sigma || !rows->queue)) {
        ts_set_error("ts_ingest_rows: null column");
        return TS_E_ARG;
    }
    FirstSeen matches(n / 2), players(n);
    std::vector<int32_t> first_row_of_match, first_row_of_player, pos(n);
    std::vector<uint32_t> count;           // per match: participants seen so far on roster 0 / 1
    std::vector<uint8_t> flags;            // per match: bit 0/1 a winner flag on roster 0/1, bit 2 an AFK participant
    int widest = 1;
    constexpr int64_t kAhead = 12;
    for (int64_t i = 0; i < n; ++i) {
        bool fresh = false;
        if (i + kAhead < n) {
            players.prefetch(rows->player[i + kAhead]);
            if (rows->match[i + kAhead] != rows->match[i + kAhead - 1]) matches.prefetch(rows->match[i + kAhead]);
        }
        // the rows of a match usually follow each other: no lookup then
        const int32_t m = (i > 0 && rows->match[i] == rows->match[i - 1]) ? row_match[i - 1] : matches.get(rows->match[i], &fresh);
        if (fresh) { match_keys[m] = rows->match[i]; first_row_of_match.push_back((int32_t)i); count.push_back(0); count.push_back(0); flags.push_back(0); }
        const int32_t p = players.get(rows->player[i], &fresh);
        if (fresh) { player_keys[p] = rows->player[i]; first_row_of_player.push_back((int32_t)i); }
        row_match[i] = m;
        row_player[i] = p;
        const int32_t t = rows->roster[i];
        pos[i] = -1;
        if (t == 0 || t == 1) {
            pos[i] = (int32_t)count[2 * m + t]++;
            if (pos[i] + 1 > widest) widest = pos[i] + 1;
            if (rows->winner[i]) flags[m] |= (uint8_t)(1u << t);
        }
        if (rows->went_afk[i] == 1) flags[m] |= 4u;
    }
    if (widest > 5) { ts_set_error("ts_ingest_rows: rosters of more than 5 participants are not supported by the kernel"); return TS_E_UNSUPPORTED; }
    const int32_t n_matches = matches.count, n_players = players.count;
    const int32_t slots = widest <= 3 ? 6 : 10, half = slots / 2;
    std::vector<uint8_t> live(n_matches);
    for (int32_t m = 0; m < n_matches; ++m) {
        const int32_t r0 = first_row_of_match[m];
        const int32_t mode = rows->game_mode[r0];
        const bool supported = mode >= 0 && mode <= 5;
        uint32_t bits = (uint32_t)(supported ? mode : 15) << TS_META_MODE_SHIFT;
        const bool invalid = rows->n_rosters[r0] != 2, afk = (flags[m] & 4u) != 0;
        if (supported && invalid) bits |= TS_META_INVALID;
        if (supported && afk) bits |= TS_META_AFK;
        live[m] = supported && !invalid && !afk;
        if (live[m]) bits |= (uint32_t)(flags[m] & 3u);          // winner flags matter for rated matches only
        meta[m] = bits;
        for (int32_t s = 0; s < slots; ++s) ids[(size_t)m * slots + s] = TS_EMPTY_SLOT;
    }
    for (int64_t i = 0; i < n; ++i) {
        row_slot[i] = pos[i] < 0 ? -1 : rows->roster[i] * half + pos[i];
        if (pos[i] >= 0 && live[row_match[i]]) ids[(size_t)row_match[i] * slots + row_slot[i]] = (uint32_t)row_player[i];
    }
    const double nan = NAN, s23 = unknown_sigma * (2.0 / 3.0);
    for (int32_t p = 0; p < n_players; ++p) {
        const int64_t r0 = first_row_of_player[p];
        double *row = table + (size_t)p * 16;
        for (int c = 0; c < 16; ++c) row[c] = nan;
        if (!isnan(rows->trueskill_mu[r0])) { row[0] = rows->trueskill_mu[r0]; row[1] = rows->trueskill_sigma[r0]; }
        const double *q = rows->queue + (size_t)r0 * 12;
        for (int c = 0; c < 6; ++c)
            if (!isnan(q[2 * c])) { row[2 + 2 * c] = q[2 * c]; row[3 + 2 * c] = q[2 * c + 1]; }
        // get_trueskill_seed (rater.py:42-62); 0 counts as absent (rater.py:45,47)
        double best = nan;
        const double ranked = rows->rank_points_ranked[r0], blitz = rows->rank_points_blitz[r0];
        if (!isnan(ranked) && ranked != 0.0) best = ranked;
        if (!isnan(blitz) && blitz != 0.0 && (isnan(best) || blitz > best)) best = blitz;
        const int32_t tier = rows->skill_tier[r0];
        if (!isnan(best)) { seed[2 * p] = best + s23; seed[2 * p + 1] = s23; }
        else if (tier >= -1 && tier <= 29) { seed[2 * p] = vst_points[tier + 1] + unknown_sigma; seed[2 * p + 1] = unknown_sigma; }
        else { seed[2 * p] = nan; seed[2 * p + 1] = nan; }       // the reference's KeyError: the first rated match fails
    }
    *n_matches_out = n_matches;
    *n_players_out = n_players;
    *slots_out = slots;
    return 0;
}
