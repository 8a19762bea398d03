/* tsrate.h -- C ABI of libtsrate.so, the B200 (sm_100a) TrueSkill rating hot path
 * of vainglorygame/analyzer.
 *
 * The reference is pure Python and has no FFI of its own; each entry point
 * below names the reference code (file:line under the analyzer tree) whose
 * work it takes over.  All pointers marked "device" are CUDA device pointers
 * owned by the caller (the Python host passes torch tensors' data_ptr()); the
 * library allocates nothing but the few hundred bytes of control words held
 * by a ts_ctx.  Every call enqueues on the caller's stream (a cudaStream_t
 * passed as void*) and returns 0 on success or a negative TS_E_* code;
 * ts_last_error() gives the text.  Nothing throws.
 */
#ifndef TSRATE_H
#define TSRATE_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define TS_MAX_SLOTS 16          /* match record width limit (teams of up to 8) */
#define TS_EMPTY_SLOT 0xFFFFFFFFu /* id of an unused slot in a match record   */
#define TS_NONE 0xFFFFFFFFu

/* match meta word (one uint32 per match) */
#define TS_META_WIN_A 1u    /* roster 0 .winner           (rater.py:144)        */
#define TS_META_WIN_B 2u    /* roster 1 .winner                                 */
#define TS_META_AFK 4u      /* some participant.went_afk  (rater.py:95-100)     */
#define TS_META_INVALID 8u  /* len(rosters) != 2          (rater.py:91-93)      */
#define TS_META_MODE_SHIFT 4 /* bits 4..7: game mode 0..5 (rater.py:70-82), 15 = unsupported */

/* per-match status byte */
enum {
    TS_ST_OK = 0,
    TS_ST_SKIP_MODE = 1,   /* unsupported game mode: untouched (rater.py:83-85)          */
    TS_ST_AFK = 2,         /* AFK gate: quality 0, no rating change (rater.py:102-106)    */
    TS_ST_INVALID = 3,     /* roster-count gate, same effect                              */
    TS_ST_FP_ERROR = 4,    /* trueskill raises FloatingPointError (w outside (0,1), zero mass) */
    TS_ST_VALUE_ERROR = 5  /* trueskill raises ValueError (sigma == 0, empty team) or a
                              player has neither a stored rating nor a seed            */
};

enum {
    TS_E_ARG = -1,      /* bad argument                                  */
    TS_E_CUDA = -2,     /* a CUDA runtime call failed                    */
    TS_E_WORKSPACE = -3,/* workspace / offsets capacity too small        */
    TS_E_UNSUPPORTED = -4
};

/* Environment of rater.py:30-37 (beta, tau, draw_probability), with the draw
 * margin calc_draw_margin() would give for a match of n players precomputed
 * by the host (n = index).  draw_is_zero = 1 selects the analytic draw limit
 * (v = -t, w = 1) for a draw (both rosters' winner flags equal) at
 * draw_probability = 0: what a 50-digit evaluation of the package's formulas lands
 * on, as far as that could be checked here -- the pinned trueskill 0.4.4 / mpmath
 * 0.19 were not available, so this branch is parity-unpinned by the reference
 * (its float backend raises FloatingPointError there).  draw_is_zero = 2 makes
 * such a match TS_ST_FP_ERROR instead, for an integrator who wants the raising
 * behaviour; 0 = draw_probability > 0 (margins in use). */
typedef struct ts_params {
    double beta;
    double tau;
    double draw_margin[TS_MAX_SLOTS + 1];
    int32_t draw_is_zero;
    int32_t reserved;
} ts_params;

/* The HBM-resident player table: n_players rows of row_pairs (mu, sigma)
 * pairs of doubles, NaN mu = SQL NULL.  Replaces the player columns
 * worker.py:184-190 loads: trueskill_mu/_sigma (shared_col) and the six
 * trueskill_<mode>_mu/_sigma pairs (mode_col[mode]).  shared_col = -1 gives
 * the single-pair ladder of BASELINE.json (one rate() per match).
 * seed (optional) holds get_trueskill_seed() per player (rater.py:42-62). */
typedef struct ts_table {
    double *data;          /* device, [n_players][row_pairs][2] */
    const double *seed;    /* device, [n_players][2] or NULL    */
    int64_t n_players;
    int32_t row_pairs;
    int32_t shared_col;
    int8_t mode_col[16];   /* column of each game mode, -1 = mode not rated */
} ts_table;

/* Result of ts_schedule, written to device memory so ts_rate can follow on
 * the same stream. */
typedef struct ts_sched_info {
    uint32_t n_rounds;     /* largest round; rounds 1..n_rounds are rated  */
    uint32_t n_gated;      /* matches in bucket 0 (they touch no player)    */
    uint32_t overflow;     /* 1 if round_offsets capacity was too small     */
    uint32_t reserved;
} ts_sched_info;

/* Optional per-match / per-participant outputs of rate_match, indexed by the
 * ORIGINAL match index (not schedule order).  Any pointer may be NULL. */
typedef struct ts_outputs {
    double *quality;     /* device [N]       match.trueskill_quality        (rater.py:103,141) */
    double *out_shared;  /* device [N][P][2] participant.trueskill_mu/sigma (rater.py:155,157) */
    double *out_delta;   /* device [N][P]    participant.trueskill_delta    (rater.py:150-153) */
    double *out_queue;   /* device [N][P][2] participant_items <mode>_mu/sigma (rater.py:167,169) */
    uint8_t *status;     /* device [N]       TS_ST_*                                          */
} ts_outputs;

typedef struct ts_ctx ts_ctx;

/* rater.py:30-37 `env = trueskill.TrueSkill(...)` -- one context per device. */
int ts_ctx_create(const ts_params *params, int device, ts_ctx **out);
void ts_ctx_destroy(ts_ctx *ctx);
const char *ts_last_error(void);
int ts_sm_count(const ts_ctx *ctx);

/* Bytes of device workspace ts_schedule needs for N matches of P slots. */
int64_t ts_schedule_workspace_bytes(int64_t n_matches, int32_t slots);

/* Replaces the ORDER BY created_at loop of worker.py:176-192: gives every
 * match the earliest conflict-free round that keeps each player's matches in
 * stream order.  ids: device [N][P] player ids in match_seq order (first P/2
 * slots roster 0, rest roster 1, TS_EMPTY_SLOT = unused); meta: device [N].
 * Outputs (device): round[N]; order[N] = match indices sorted by (round,
 * match_seq); round_offsets[offsets_cap] with bucket r at
 * [round_offsets[r], round_offsets[r+1]), r = 0 (gated) .. n_rounds;
 * ids_sorted[N][P] / meta_sorted[N] = the records permuted into that order;
 * info.  With n_rounds_host == NULL the call only enqueues work: the number of
 * rounds stays on the device (the final sort sizes itself there), and if
 * round_offsets is too small for it, info->overflow is set, ts_rate on that
 * schedule rates nothing and counts one value error.  With n_rounds_host the
 * call synchronises the stream once, stores n_rounds there and returns
 * TS_E_WORKSPACE on overflow. */
int ts_schedule(ts_ctx *ctx, const uint32_t *ids, const uint32_t *meta, int64_t n_matches,
                int32_t slots, int64_t n_players, const int8_t mode_col[16],
                uint32_t *round, uint32_t *order, uint32_t *round_offsets, int64_t offsets_cap,
                uint32_t *ids_sorted, uint32_t *meta_sorted, ts_sched_info *info,
                void *workspace, int64_t workspace_bytes, uint32_t *n_rounds_host, void *stream);

/* ts_schedule without any allocation.  No counterpart in the reference: it is
 * for hosts that embed this library where device memory is not theirs to
 * allocate -- a plugin inside a service that hands out device memory from its
 * own pool, a process whose HBM is already reserved by another framework -- and
 * that therefore can call neither cudaMalloc (for a workspace) nor
 * ts_ctx_create (which allocates its control words).  Inputs and outputs are
 * still the caller's; the scratch is a 2 MiB array built into the library image
 * on every GPU that loads it.  Same arguments and results as ts_schedule, minus
 * ctx / workspace, plus the device ordinal.
 *   - n_matches * slots must not exceed TS_STATIC_MAX_SLOT_ENTRIES, else
 *     TS_E_WORKSPACE (ts_last_error says so);
 *   - the first call on a device resolves the array's device address and caches
 *     it: make that call before any stream capture, and from one host thread at a
 *     time per device (the cache is filled under a lock, but two first calls would
 *     also share the scratch, see below);
 *   - like ts_schedule it synchronises `stream` once when n_rounds_host is given,
 *     and only enqueues work when it is NULL;
 *   - all calls share the array; it is not cleared between calls except for what
 *     the scheduler needs zeroed, which each call clears on `stream` itself;
 *     calls whose GPU work can overlap race -- serialise them;
 *   - performs no cudaMalloc / cudaFree / allocating library call (the test
 *     checks that the free device memory is the same before and after). */
#define TS_STATIC_MAX_SLOT_ENTRIES 65536
int ts_schedule_static(int device, const uint32_t *ids, const uint32_t *meta, int64_t n_matches,
                       int32_t slots, int64_t n_players, const int8_t mode_col[16],
                       uint32_t *round, uint32_t *order, uint32_t *round_offsets, int64_t offsets_cap,
                       uint32_t *ids_sorted, uint32_t *meta_sorted, ts_sched_info *info,
                       uint32_t *n_rounds_host, void *stream);
int64_t ts_schedule_static_max_slot_entries(void);

/* Replaces rater.py:108-169 for every match of the batch: gather the players'
 * pairs, quality, rate(shared), delta, rate(queue), scatter -- one persistent
 * kernel walking the rounds of a schedule.  lanes_per_match: 0 = default, or
 * one of 1,2,4,8,16,32 (how many lanes of a warp share one match). */
int ts_rate(ts_ctx *ctx, const ts_table *table, const uint32_t *ids_sorted,
            const uint32_t *meta_sorted, const uint32_t *order, const uint32_t *round_offsets,
            const ts_sched_info *info, int64_t n_matches, int32_t slots,
            const ts_outputs *outputs, int32_t lanes_per_match, void *stream);

/* Streaming (dataflow) form of the same path: no rounds, no barriers.  Tables:
 * a single-pair ladder (row_pairs == 1, shared_col == -1), or the reference's
 * shared pair plus queue columns (shared_col >= 0; quality, shared / delta /
 * queue outputs and the NULL fallbacks of rater.py:115-130 are all produced).
 * With ONE rated queue column a player's row is 64 bytes under one version;
 * with several (a batch that mixes game modes, worker.py:176-192) every queue
 * pair is versioned by its own ordinal and the ordinals travel with the shared
 * pair (csrc/stream_pair.cu) -- the links are the same single pred array.
 * ts_link_pred does the first half of the scheduler only -- radix sort by
 * player id and chain linking -- and writes pred[N][P]: for every slot the
 * stream-wide number of that player's previous rated match (TS_NONE if none,
 * and for empty slots and gated matches).  ts_rate_stream then rates
 * the matches in stream order (worker.py:192), each match waiting until its
 * players' rows carry the version pred names.  Results are identical to
 * ts_schedule + ts_rate.  ids / meta are in ORIGINAL order; outputs too.
 * workspace: ts_rate_stream_table_workspace_bytes(table) -- 32 bytes per player
 * for a single-pair ladder, 64 for shared + one queue, 64 + 32 per rated queue
 * column otherwise; ts_rate_stream_workspace_bytes(U) = 64 U covers the first two.
 *
 * Windows (worker.py:18-20 micro-batches; ladders beyond 2^32 slot entries):
 * a stream is processed window after window by passing seq_base = number of
 * matches in the earlier windows, and `last` = device uint32[n_players]
 * holding every player's latest match so far (all TS_NONE before the first
 * window; ts_link_pred reads it for chain heads and then updates it).  The
 * versioned rows stay in `workspace` between windows: pass TS_STREAM_LOAD with
 * the first window (rows are built from the table) and TS_STREAM_STORE with
 * the last (table is written back); a one-shot call passes both, seq_base 0
 * and last = NULL.  Stream-wide match numbers must stay below 2^32 - 1. */
#define TS_STREAM_LOAD 1
#define TS_STREAM_STORE 2
int64_t ts_link_workspace_bytes(int64_t n_matches, int32_t slots);
int ts_link_pred(ts_ctx *ctx, const uint32_t *ids, const uint32_t *meta, int64_t n_matches, int32_t slots,
                 int64_t n_players, const int8_t mode_col[16], int64_t seq_base, uint32_t *last, uint32_t *pred,
                 void *workspace, int64_t workspace_bytes, void *stream);
int64_t ts_rate_stream_workspace_bytes(int64_t n_players);
int64_t ts_rate_stream_table_workspace_bytes(const ts_table *table);
/* 1 if the last ts_rate_stream / ts_rate_stream_sharded on this context gave up
 * because some match's links were never satisfied (pred does not belong to the
 * stream it was given with): the table is then unusable.  A watchdog inside the
 * kernel, so that bad links cannot hang the GPU.  Synchronises the stream. */
int ts_stream_stalled(ts_ctx *ctx, void *stream);
int ts_rate_stream(ts_ctx *ctx, const ts_table *table, const uint32_t *ids, const uint32_t *meta,
                   const uint32_t *pred, int64_t n_matches, int32_t slots, int64_t seq_base, int32_t flags,
                   const ts_outputs *outputs, int32_t lanes_per_match, void *workspace, int64_t workspace_bytes,
                   void *stream);

/* The streaming form for ladders whose rows do not fit L2 (single-pair ladders only): a row
 * lives where it will be READ -- rows[m * slots + s] is the rating the player in slot s of match
 * m brings to that match -- so a match reads its rows as one contiguous piece and only the
 * writes scatter.  ts_link_succ sorts like ts_link_pred and writes succ[N][P] (the slot index
 * m' * slots + s' of the same player's next rated match, TS_NONE after the last) and every
 * player's first row (taken from the table); ts_rate_slots rates in stream order, a match waiting
 * until its own rows carry `epoch`, and stores every new row at succ, or into the table after the
 * player's last match.  Results are bit-identical to ts_rate_stream.  rows: ts_slot_rows_bytes
 * (32 bytes per slot of the window), 32-byte aligned, zeroed once when allocated; `epoch` (not 0,
 * not TS_NONE) must differ from the one of the buffer's previous use -- it is what validates a row,
 * so the buffer is never cleared.  Windows need no carried state: the table holds it between them.
 * A player listed twice in one match is not supported (ts_shard_status reports TS_SHARD_DUPLICATE).
 * No counterpart in the reference. */
int64_t ts_slot_rows_bytes(int64_t n_matches, int32_t slots);
int ts_link_succ(ts_ctx *ctx, const uint32_t *ids, const uint32_t *meta, int64_t n_matches, int32_t slots,
                 const ts_table *table, uint32_t epoch, uint32_t *succ, void *rows, void *workspace,
                 int64_t workspace_bytes, void *stream);
int ts_rate_slots(ts_ctx *ctx, const ts_table *table, const uint32_t *ids, const uint32_t *meta, const uint32_t *succ,
                  int64_t n_matches, int32_t slots, uint32_t epoch, const ts_outputs *outputs, int32_t lanes_per_match,
                  void *rows, void *stream);

/* One ladder over the G GPUs of a node (BASELINE.json configs[4]), one
 * process per GPU, G a power of two <= 16.  No counterpart in the reference (single
 * process, single thread: worker.py:176-192 is the one sequential order this reproduces).
 *
 * Matches: blocks of TS_SHARD_BLOCK consecutive matches go round-robin to the GPUs --
 * match m of a window lives on rank (m / 32) % G at local index ((m / 32) / G) * 32 +
 * m % 32 -- and every rank holds only its own slice of ids / meta.
 * Table: player p is owned by rank p % G (local row p / G): the owner holds the
 * player's (mu, sigma) before the first and after the last match of a window.
 * Rows travel to where they are needed: every rank has a mailbox of n_players
 * versioned rows (32 bytes each); the rank that rates a player's match sends the new
 * row, tagged with the number of the player's NEXT match, to the rank that will rate
 * that match, and a match polls only its own GPU's mailbox until every row carries the
 * match's own number.  A row for another GPU is a 32-byte message appended to the
 * sender's ring in the receiver's memory (consecutive addresses: scattered stores over
 * several peers' memory collapse on B200), and router blocks of the receiver's own grid
 * put it into its mailbox row.  After a player's last match the pair goes back to the
 * owner's table shard the same way.  No inter-GPU barrier and no collective while
 * rating.  ts_shard_reset clears mailbox, rings and counts before a ladder's first window;
 * every window's messages are appended behind the earlier ones (size ring_cap for the run).
 * Links: ts_shard_link takes the slots of the players it owns -- sent to it by the ranks that
 * hold the matches (ts_shard_count / ts_shard_send below: every rank reads its own slice only),
 * or, without those two calls, by reading the whole id stream once itself (its peers' slices
 * over NVLink) -- sorts them, and sends succ[m][s] (the
 * stream-wide number of that player's next rated match, TS_NONE if none) to the rank
 * that rates m -- as messages appended to that rank's inbox in large coalesced pieces,
 * which ts_shard_rate scatters into the succ slice locally before it rates (scattered
 * 4-byte stores into a peer's memory are partial-sector writes and crawl) -- together
 * with every owned player's first row.  TS_SHARD_OVERFLOW: workspace or inbox too small.
 * A player listed twice in one match is not supported here (TS_SHARD_DUPLICATE).
 *
 * All arrays of a ts_shard_view are device pointers valid on THIS GPU: index rank =
 * own allocations, the others peer mappings (ts_shard_alloc / ts_shard_open).
 * Call order per window, with a barrier across the ranks at every "|":
 *   slices and table shards in place, ts_shard_reset | [ts_shard_count | ts_shard_send |] ts_shard_link | ts_shard_rate |
 * Windows (seq_base > 0) continue a ladder: seq_base must be a multiple of 32 * G and
 * stream-wide match numbers stay below 2^32 - 1; ts_shard_reset is for the first only. */
#define TS_SHARD_BLOCK 32
#define TS_SHARD_MAX 16
typedef struct ts_shard_view {
    int32_t rank, n_shards;
    const uint32_t *ids[TS_SHARD_MAX];       /* [n_local(g)][slots]                              */
    const uint32_t *meta[TS_SHARD_MAX];      /* [n_local(g)]                                     */
    uint32_t *succ[TS_SHARD_MAX];            /* [n_local(g)][slots], written by the players' owners */
    unsigned long long *rows[TS_SHARD_MAX];  /* mailbox: n_players rows of four 64-bit words      */
    double *table[TS_SHARD_MAX];             /* [ceil((n_players - g) / G)][2]                    */
    unsigned long long *inbox[TS_SHARD_MAX]; /* link messages for rank g: G * inbox_cap of 8 bytes */
    uint32_t *inbox_count[TS_SHARD_MAX];     /* 256 words: messages per (range of slots, sender)  */
    int64_t inbox_cap;                       /* per sender: about n_matches * slots / G^2, with room for skew */
    unsigned long long *ring[TS_SHARD_MAX];  /* row messages for rank g: [G senders][ring_cap] of 32 bytes   */
    unsigned long long *ring_final[TS_SHARD_MAX]; /* [16]: every sender's final message count for rank g      */
    uint32_t *ring_tail[TS_SHARD_MAX];       /* [48], own rank only: places reserved, blocks done, consumed   */
    int64_t ring_cap;                        /* per sender: about (n_matches * slots + n_players) / G^2 + skew */
    int64_t ring_epoch;                      /* 1..255, the same on all ranks: +1 for every ts_shard_reset, 1 again after 255 */
    /* link exchange (ts_shard_count / ts_shard_send); all NULL / 0: ts_shard_link reads the whole stream itself */
    uint16_t *link_counts[TS_SHARD_MAX];     /* [G senders][G owners][link_blocks_cap]: owned live slots per block of 32 matches */
    uint32_t *in_keys[TS_SHARD_MAX];         /* [in_cap] local index of the owned player of every pair sent to rank g, stream order */
    uint32_t *in_vals[TS_SHARD_MAX];         /* [in_cap] its slot index m * slots + s in the window                              */
    uint32_t *in_count[TS_SHARD_MAX];        /* [1] pairs rank g received (own rank only)                                        */
    int64_t in_cap;                          /* pairs a rank can receive: about n_matches * slots / G, with room for skew        */
    int64_t link_blocks_cap;                 /* a multiple of 8, >= ceil(ceil(n_matches / 32) / G)                               */
} ts_shard_view;
/* ts_shard_status bits */
#define TS_SHARD_STALLED 1    /* a match's rows never arrived (watchdog): tables unusable      */
#define TS_SHARD_OVERFLOW 2   /* ts_shard_link: workspace too small for the owned slots        */
#define TS_SHARD_DUPLICATE 4  /* a player is listed twice in one match                          */
int64_t ts_shard_local_matches(int64_t n_matches, int32_t rank, int32_t n_shards);
/* inbox_cap to allocate for windows of up to n_matches (8 bytes per message, G senders per inbox) */
int64_t ts_shard_inbox_cap(int64_t n_matches, int32_t slots, int32_t n_shards);
/* workspace for ts_shard_link on a window of n_slot_entries = n_matches * slots: 8 bytes per slot
 * entry (staging of the compaction) plus the sort's buffers for `owned_cap` (player, slot) pairs;
 * a rank owns about n_slot_entries / n_shards of them (all of them in the worst case) */
int64_t ts_shard_link_workspace_bytes(int64_t n_slot_entries, int64_t owned_cap);
int ts_shard_alloc(int device, int64_t bytes, void **ptr, void *handle64);
int ts_shard_open(int device, const void *handle64, void **ptr);
int ts_shard_close(void *ptr);
int ts_shard_free(void *ptr);
int ts_shard_reset(ts_ctx *ctx, const ts_shard_view *view, int64_t n_players, void *stream);
int ts_shard_link(ts_ctx *ctx, const ts_shard_view *view, int64_t n_matches, int32_t slots, int64_t n_players,
                  int8_t mode, int64_t seq_base, void *workspace, int64_t workspace_bytes, void *stream);
/* Link exchange: the same links with every rank reading only its OWN slice of the stream (link
 * work per rank ~ 1/G of the window instead of all of it).  ts_shard_count counts, per block of 32
 * matches of the own slice and per owner, the live slots of that owner's players and copies the
 * counts to every rank's link_counts; after a barrier ts_shard_send scans the counts of all ranks in
 * stream order and stores every (local player index, slot index) pair at its stream-ordered place
 * in its owner's in_keys / in_vals (consecutive pairs of a block to consecutive addresses); after
 * another barrier ts_shard_link, finding in_keys set in the view, sorts those pairs instead of
 * reading the stream (pass n_slot_entries = 0 to ts_shard_link_workspace_bytes then).  Call order:
 *   ts_shard_count | ts_shard_send | ts_shard_link | ts_shard_rate |
 * TS_SHARD_OVERFLOW: a rank was sent more than in_cap pairs.  Results are identical to the
 * links ts_shard_link makes on its own. */
int64_t ts_shard_exchange_workspace_bytes(int64_t n_matches, int32_t n_shards);
int ts_shard_count(ts_ctx *ctx, const ts_shard_view *view, int64_t n_matches, int32_t slots, int64_t n_players,
                   int8_t mode, void *stream);
int ts_shard_send(ts_ctx *ctx, const ts_shard_view *view, int64_t n_matches, int32_t slots, int64_t n_players,
                  int8_t mode, void *workspace, int64_t workspace_bytes, void *stream);
/* outputs: indexed by the LOCAL match index of this rank.  emulate_all = 1: this one
 * launch plays every rank of the view on one GPU (all pointers local; testing and
 * single-GPU runs of the sharded code path); outputs are then indexed by the window's
 * match number. */
int ts_shard_rate(ts_ctx *ctx, const ts_shard_view *view, int64_t n_matches, int32_t slots, int64_t n_players,
                  int8_t mode, int64_t seq_base, const ts_outputs *outputs, int32_t lanes_per_match,
                  int32_t emulate_all, void *stream);
/* TS_SHARD_* bits raised on this context since the last ts_shard_reset.  Synchronises the stream. */
int ts_shard_status(ts_ctx *ctx, void *stream);

/* Number of matches that ended in TS_ST_FP_ERROR / TS_ST_VALUE_ERROR since the
 * last ts_reset_errors (the reference raises and fails the whole batch,
 * worker.py:108-120,195-197).  Synchronises the stream. */
int64_t ts_error_count(ts_ctx *ctx, void *stream);
/* the two kinds apart, so that a host can raise what the reference would raise without
 * asking for the per-match status bytes */
int ts_error_counts(ts_ctx *ctx, int64_t *fp_errors, int64_t *value_errors, void *stream);
/* also clears the stall flag of the streaming kernels, which is sticky: after a stall
 * every streaming launch on this context returns at once until this call */
int ts_reset_errors(ts_ctx *ctx, void *stream);

/* get_trueskill_seed (rater.py:42-62) for a column of players, on the device:
 * rank points as doubles (NaN = None; 0 = absent, rater.py:45,47), skill_tier
 * int8 (-1..29; anything else with no rank points is the reference's KeyError:
 * seed stays NaN and bad[i] = 1).  vst_points[tier + 1] = rater.py:13-27.
 * Any input column may be NULL (= all None); bad may be NULL. */
int ts_seed_table(const double *rank_points_ranked, const double *rank_points_blitz, const int8_t *skill_tier,
                  int64_t n_players, double unknown_player_sigma, const double vst_points[31], double *seed,
                  uint8_t *bad, void *stream);

/* Ingest helper: the meta word uses its low 8 bits only, so a host may ship the
 * column as bytes (a quarter of its PCIe volume) and widen it on the device.
 * No counterpart in the reference (it has no columnar format). */
int ts_widen_meta(const uint8_t *meta8, int64_t n_matches, uint32_t *meta, void *stream);
/* The id column at the width it needs: `count` ids packed as a little-endian bit stream of
 * `bits` bits each (bits = ceil(log2(n_players + 1)); the all-ones value is TS_EMPTY_SLOT),
 * widened to the uint32 the kernels read.  worker.py:176-191 loads nothing wider than a
 * player's identity per participant; for a million players this is 20/32 of the PCIe volume
 * of the column that bounds the end-to-end call.  No counterpart in the reference. */
int ts_unpack_ids(const uint8_t *packed, int32_t bits, int64_t count, uint32_t *ids, void *stream);

/* Columnar ingest on the host (no device work): the participant rows the query of
 * worker.py:176-191 yields -- one row per participant, matches in created_at order -- become the
 * columns above.  game_mode 0..5 in rater.py:70-82's order (anything else: unsupported), roster 0 / 1,
 * NaN = None / NULL for the doubles, queue = the player's six trueskill_<mode> pairs [n_rows][6][2].
 * Outputs, all caller-allocated for n_rows entries (ids: 10 * n_rows, table: 16 * n_rows doubles as
 * [players][8][2] with col 0 shared and 1..6 the queues, seed: 2 * n_rows): matches and players in
 * order of first appearance with their keys, every row's match / slot (-1: not on roster 0 or 1) /
 * player index, ids [n_matches][slots] (slots = 6 or 10 by the widest roster), meta, table, seed
 * (get_trueskill_seed, rater.py:42-62; NaN where the reference raises KeyError). */
typedef struct ts_row_columns {
    int64_t n_rows;
    const int64_t *match, *player;
    const int32_t *game_mode, *n_rosters, *roster, *skill_tier;
    const uint8_t *winner, *went_afk;
    const double *rank_points_ranked, *rank_points_blitz, *trueskill_mu, *trueskill_sigma, *queue;
} ts_row_columns;
int ts_ingest_rows(const ts_row_columns *rows, double unknown_player_sigma, const double vst_points[31],
                   int64_t *match_keys, int64_t *player_keys, int32_t *row_match, int32_t *row_slot, int32_t *row_player,
                   uint32_t *ids, uint32_t *meta, double *table, double *seed, int64_t *n_matches, int64_t *n_players,
                   int32_t *slots);

/* Synthetic ladders of BASELINE.json's configs (SURVEY.md 8d "Generator"):
 * counter-based, so any window [first, first+count) can be produced on its
 * own.  No counterpart in the reference (it has no benchmark input). */
typedef struct ts_gen {
    uint64_t seed;
    uint64_t n_players;
    int32_t slots;
    int32_t mode;
    uint32_t draw_thr;   /* P(draw) * 2^32     */
    uint32_t afk_thr;    /* P(AFK gate) * 2^32 */
    int32_t zipf;        /* 0 uniform, 1 Zipf-Mandelbrot (rank + q)^-0.8 */
    double zipf_a;       /* q^0.2               */
    double zipf_b;       /* (U + q)^0.2 - q^0.2 */
    double zipf_q;
} ts_gen;
int ts_gen_matches(const ts_gen *gen, int64_t first, int64_t count, uint32_t *ids, uint32_t *meta,
                   void *stream);
int ts_gen_table(uint64_t seed, int64_t first, int64_t count, double *table, void *stream);

#ifdef __cplusplus
}
#endif
#endif /* TSRATE_H */
