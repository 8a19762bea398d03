/* ORACLE O2 -- test infrastructure, NOT product code.
 *
 * Sequential CPU restatement of the analyzer's rating hot path in columnar
 * form, used only by tests/, __graft_entry__.smoke() and bench.py's
 * cpu_baseline leg as the checker for the CUDA path.
 *
 * What it follows (file:line under /root/reference):
 *   worker.py:176-192   one pass over the matches in created_at order
 *   rater.py:70-85      mode -> rating column, unsupported mode = skip
 *   rater.py:90-106     roster-count / AFK gate: quality = 0, no rating change
 *   rater.py:108-136    per participant: shared pair (stored, else seed),
 *                       queue pair (stored, else the shared pair just read)
 *   rater.py:141        quality on the queue-specific pre-update ratings
 *   rater.py:144-157    rate(shared) ; trueskill_delta ; write-back
 *   rater.py:161-169    rate(queue)  ; write-back
 * The arithmetic of rate()/quality() is the third-party trueskill==0.4.4
 * package (requirements.txt:6; not in /root/reference, not installable here).
 * For two teams its factor graph collapses to the closed form in SURVEY.md
 * Appendix A, evaluated here in __float128 (libquadmath) -- or in fp64 for the
 * "strong CPU baseline" -- and rounded to double after every match exactly
 * where the package calls float().  tests/test_oracle_c.py checks this file
 * against the factor-graph shim (oracle/shim) at 50 digits.
 *
 * Parity status: pinned on the public TrueSkill known-answer vectors through
 * the shim; the reference's own tests pin no rate()/quality() value
 * ("parity unpinned" by the reference).
 *
 * Also here: the schedule oracle (one sequential integer pass giving every
 * match its ASAP round) and the host twin of the synthetic generator, which
 * must be bit-identical to the CUDA generator in analyzer_b200/csrc.
 */
#include <math.h>
#include <quadmath.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

#define TSO_EMPTY 0xFFFFFFFFu
#define TSO_MAXP 16
#define META_WIN_A 1u
#define META_WIN_B 2u
#define META_AFK 4u
#define META_INVALID 8u
#define META_MODE(m) (((m) >> 4) & 15u)

enum { ST_OK = 0, ST_SKIP_MODE = 1, ST_AFK = 2, ST_INVALID = 3, ST_FP_ERROR = 4, ST_VALUE_ERROR = 5 };

typedef struct {
    double beta;
    double tau;
    double draw_margin[TSO_MAXP + 1]; /* indexed by players in the match */
    int32_t draw_is_zero;             /* draw_probability == 0: analytic limit v=-t, w=1 */
} tso_params;

/* ------------------------------------------------------------------ */
/* closed-form two-team update, generic over the float type            */
/* ------------------------------------------------------------------ */
#define DEFINE_UPDATE(NAME, T, SQRT, EXP, ERFC, FABS, LIT)                                      \
    static inline T NAME##_pdf(T x) { return EXP(-x * x / LIT(2.0)) / SQRT(LIT(2.0) * LIT(3.14159265358979323846264338327950288419716939937510)); } \
    static inline T NAME##_cdf(T x) { return ERFC(-x / SQRT(LIT(2.0))) / LIT(2.0); }            \
    /* returns 0 ok, ST_FP_ERROR, ST_VALUE_ERROR. first team = winner (or roster 0 in a draw) */ \
    static int NAME(const tso_params *pr, int n, const double *mu, const double *sg,           \
                    const int *team /*0 = first, 1 = second*/, int is_draw, double *mu_out,     \
                    double *sg_out)                                                             \
    {                                                                                           \
        T tau2 = (T)pr->tau * (T)pr->tau, beta2 = (T)pr->beta * (T)pr->beta;                    \
        T s[TSO_MAXP], c2 = 0, dmu = 0;                                                         \
        int n0 = 0, n1 = 0;                                                                     \
        for (int i = 0; i < n; i++) {                                                           \
            if (sg[i] == 0.0) return ST_VALUE_ERROR;                                            \
            s[i] = (T)sg[i] * (T)sg[i] + tau2;                                                  \
            c2 += s[i] + beta2;                                                                 \
            if (team[i] == 0) { dmu += (T)mu[i]; n0++; } else { dmu -= (T)mu[i]; n1++; }        \
        }                                                                                       \
        if (n0 == 0 || n1 == 0) return ST_VALUE_ERROR;                                          \
        T c = SQRT(c2);                                                                         \
        T t = dmu / c;                                                                          \
        T eps = (T)pr->draw_margin[n] / c;                                                      \
        T v, w;                                                                                 \
        if (!is_draw) {                                                                         \
            T x = t - eps;                                                                      \
            T mass = NAME##_cdf(x);                                                             \
            v = (mass != 0) ? NAME##_pdf(x) / mass : -x;                                        \
            w = v * (v + x);                                                                    \
            if (!(w > 0 && w < 1)) return ST_FP_ERROR;                                          \
        } else if (pr->draw_is_zero) {                                                          \
            v = -t; w = 1;                                                                      \
        } else {                                                                                \
            T at = FABS(t);                                                                     \
            T a = eps - at, b = -eps - at;                                                      \
            T mass = NAME##_cdf(a) - NAME##_cdf(b);                                             \
            if (mass == 0) return ST_FP_ERROR;                                                  \
            T vv = (NAME##_pdf(b) - NAME##_pdf(a)) / mass;                                      \
            w = vv * vv + (a * NAME##_pdf(a) - b * NAME##_pdf(b)) / mass;                       \
            v = (t < 0) ? -vv : vv;                                                             \
        }                                                                                       \
        for (int i = 0; i < n; i++) {                                                           \
            T g = s[i] / c;                                                                     \
            T m2 = (team[i] == 0) ? (T)mu[i] + g * v : (T)mu[i] - g * v;                        \
            T var = s[i] * (1 - (s[i] / c2) * w);                                               \
            mu_out[i] = (double)m2;                                                             \
            sg_out[i] = (double)SQRT(var);                                                      \
        }                                                                                       \
        return ST_OK;                                                                           \
    }

#define LITQ(x) x##Q
#define LITD(x) x
DEFINE_UPDATE(upd_q, __float128, sqrtq, expq, erfcq, fabsq, LITQ)
DEFINE_UPDATE(upd_d, double, sqrt, exp, erfc, fabs, LITD)

/* quality(): rater.py:141, trueskill's matrix expression for two teams, in
 * plain doubles like the package (it uses math.exp / math.sqrt on floats). */
static double quality2(const tso_params *pr, int n, const double *mu, const double *sg, const int *team)
{
    double b2n = (double)n * pr->beta * pr->beta, ss = 0, d = 0;
    for (int i = 0; i < n; i++) {
        ss += sg[i] * sg[i];
        d += team[i] == 0 ? mu[i] : -mu[i];
    }
    double den = b2n + ss;
    return exp(-0.5 * d * d / den) * sqrt(b2n / den);
}

/* One match against the table.  Returns the status; writes nothing on error. */
static int rate_one(const tso_params *pr, int P, const uint32_t *ids, uint32_t meta,
                    double *table, int C, int shared_col, const int8_t *mode_col,
                    const double *seed, int use_quad,
                    double *quality, double *o_shared, double *o_delta, double *o_queue)
{
    uint32_t mode = META_MODE(meta);
    int qcol = (mode < 6) ? mode_col[mode] : -1;
    if (qcol < 0) return ST_SKIP_MODE;
    if (meta & META_INVALID) { if (quality) *quality = 0; return ST_INVALID; }
    if (meta & META_AFK) { if (quality) *quality = 0; return ST_AFK; }

    int half = P / 2, n = 0;
    int slot[TSO_MAXP], team[TSO_MAXP], had_shared[TSO_MAXP];
    double smu[TSO_MAXP], ssg[TSO_MAXP], qmu[TSO_MAXP], qsg[TSO_MAXP];
    /* winner first (stable): rank = !winner, rater.py:144 */
    int rank0 = !(meta & META_WIN_A), rank1 = !(meta & META_WIN_B);
    int first = (rank1 < rank0) ? 1 : 0;
    int is_draw = rank0 == rank1;
    for (int s = 0; s < P; s++) {
        uint32_t p = ids[s];
        if (p == TSO_EMPTY) continue;
        double *row = table + (size_t)p * C * 2;
        int r = s / half;
        slot[n] = s;
        team[n] = (r == first) ? 0 : 1;
        double m_s = NAN, s_s = NAN;
        had_shared[n] = 0;
        if (shared_col >= 0) {
            m_s = row[shared_col * 2]; s_s = row[shared_col * 2 + 1];
            had_shared[n] = !isnan(m_s);
            if (isnan(m_s)) {                     /* rater.py:118-119 */
                if (!seed) return ST_VALUE_ERROR;
                m_s = seed[(size_t)p * 2]; s_s = seed[(size_t)p * 2 + 1];
                if (isnan(m_s)) return ST_VALUE_ERROR;
            }
        }
        smu[n] = m_s; ssg[n] = s_s;
        double m_q = row[qcol * 2], s_q = row[qcol * 2 + 1];
        if (isnan(m_q)) {                         /* rater.py:127-130 */
            if (shared_col >= 0) { m_q = m_s; s_q = s_s; }
            else {
                if (!seed) return ST_VALUE_ERROR;
                m_q = seed[(size_t)p * 2]; s_q = seed[(size_t)p * 2 + 1];
                if (isnan(m_q)) return ST_VALUE_ERROR;
            }
        }
        qmu[n] = m_q; qsg[n] = s_q;
        n++;
    }
    double nsm[TSO_MAXP], nss[TSO_MAXP], nqm[TSO_MAXP], nqs[TSO_MAXP];
    int st;
    if (n == 0) return ST_VALUE_ERROR;
    for (int i = 0; i < n; i++) if (qsg[i] == 0.0) return ST_VALUE_ERROR;
    double q = quality2(pr, n, qmu, qsg, team);
    if (shared_col >= 0) {
        st = use_quad ? upd_q(pr, n, smu, ssg, team, is_draw, nsm, nss)
                      : upd_d(pr, n, smu, ssg, team, is_draw, nsm, nss);
        if (st) return st;
    }
    st = use_quad ? upd_q(pr, n, qmu, qsg, team, is_draw, nqm, nqs)
                  : upd_d(pr, n, qmu, qsg, team, is_draw, nqm, nqs);
    if (st) return st;
    if (quality) *quality = q;
    for (int i = 0; i < n; i++) {
        double *row = table + (size_t)ids[slot[i]] * C * 2;
        if (shared_col >= 0) {
            row[shared_col * 2] = nsm[i]; row[shared_col * 2 + 1] = nss[i];
            if (o_shared) { o_shared[slot[i] * 2] = nsm[i]; o_shared[slot[i] * 2 + 1] = nss[i]; }
            if (o_delta) o_delta[slot[i]] = had_shared[i] ? (nsm[i] - nss[i]) - (smu[i] - ssg[i]) : 0.0;
        }
        row[qcol * 2] = nqm[i]; row[qcol * 2 + 1] = nqs[i];
        if (o_queue) { o_queue[slot[i] * 2] = nqm[i]; o_queue[slot[i] * 2 + 1] = nqs[i]; }
    }
    return ST_OK;
}

/* Sequential pass in match_seq order (worker.py:192).  Optional outputs may be
 * NULL.  Returns the number of matches whose status is an error (>= 4). */
int64_t tso_rate_sequential(const tso_params *pr, int64_t N, int P, const uint32_t *ids,
                            const uint32_t *meta, double *table, int C, int shared_col,
                            const int8_t *mode_col, const double *seed, int use_quad,
                            double *quality, double *o_shared, double *o_delta, double *o_queue,
                            uint8_t *status)
{
    int64_t bad = 0;
    if (P > TSO_MAXP || (P & 1)) return -1;
    for (int64_t m = 0; m < N; m++) {
        int st = rate_one(pr, P, ids + m * P, meta[m], table, C, shared_col, mode_col, seed, use_quad,
                          quality ? quality + m : NULL, o_shared ? o_shared + m * P * 2 : NULL,
                          o_delta ? o_delta + m * P : NULL, o_queue ? o_queue + m * P * 2 : NULL);
        if (status) status[m] = (uint8_t)st;
        if (st >= ST_FP_ERROR) bad++;
    }
    return bad;
}

/* Same arithmetic (fp64), executed round by round over a given schedule with
 * OpenMP: the best-effort multi-core CPU baseline. */
int64_t tso_rate_rounds_omp(const tso_params *pr, int64_t N, int P, const uint32_t *ids,
                            const uint32_t *meta, double *table, int C, int shared_col,
                            const int8_t *mode_col, const double *seed, const uint32_t *order,
                            const uint32_t *round_offsets, uint32_t n_rounds, int threads)
{
    int64_t bad = 0;
    (void)N;
#ifdef _OPENMP
    if (threads > 0) omp_set_num_threads(threads);
#endif
    for (uint32_t r = 1; r <= n_rounds; r++) {
        int64_t lo = round_offsets[r], hi = round_offsets[r + 1];
#pragma omp parallel for schedule(static) reduction(+ : bad)
        for (int64_t j = lo; j < hi; j++) {
            int64_t m = order[j];
            int st = rate_one(pr, P, ids + m * P, meta[m], table, C, shared_col, mode_col, seed, 0,
                              NULL, NULL, NULL, NULL);
            if (st >= ST_FP_ERROR) bad++;
        }
    }
    return bad;
}

int tso_max_threads(void)
{
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}

/* ------------------------------------------------------------------ */
/* schedule oracle: ASAP round of every match, one sequential pass     */
/* ------------------------------------------------------------------ */
static int is_gated(uint32_t meta, const int8_t *mode_col)
{
    uint32_t mode = META_MODE(meta);
    if (mode >= 6 || mode_col[mode] < 0) return 1;
    return (meta & (META_AFK | META_INVALID)) != 0;
}

/* round[m] = 0 for gated matches (they touch no player: rater.py:83-85,102-106),
 * else 1 + max over its players of the round of that player's previous rated
 * match.  order = matches sorted by (round, match_seq); round_offsets has
 * n_rounds + 2 entries: bucket r is [round_offsets[r], round_offsets[r+1]).
 * Returns n_rounds (largest round), or -1 if offsets_cap is too small. */
int64_t tso_schedule(int64_t N, int P, const uint32_t *ids, const uint32_t *meta,
                     const int8_t *mode_col, int64_t U, uint32_t *round, uint32_t *order,
                     uint32_t *round_offsets, int64_t offsets_cap)
{
    uint32_t *last = (uint32_t *)calloc((size_t)U, sizeof(uint32_t));
    uint32_t R = 0;
    for (int64_t m = 0; m < N; m++) {
        if (is_gated(meta[m], mode_col)) { round[m] = 0; continue; }
        uint32_t r = 0;
        for (int s = 0; s < P; s++) {
            uint32_t p = ids[m * P + s];
            if (p != TSO_EMPTY && last[p] > r) r = last[p];
        }
        r += 1;
        for (int s = 0; s < P; s++) {
            uint32_t p = ids[m * P + s];
            if (p != TSO_EMPTY) last[p] = r;
        }
        round[m] = r;
        if (r > R) R = r;
    }
    free(last);
    if ((int64_t)R + 2 > offsets_cap) return -1;
    memset(round_offsets, 0, ((size_t)R + 2) * sizeof(uint32_t));
    for (int64_t m = 0; m < N; m++) round_offsets[round[m] + 1]++;
    for (uint32_t r = 0; r <= R; r++) round_offsets[r + 1] += round_offsets[r];
    uint32_t *cur = (uint32_t *)malloc(((size_t)R + 1) * sizeof(uint32_t));
    memcpy(cur, round_offsets, ((size_t)R + 1) * sizeof(uint32_t));
    for (int64_t m = 0; m < N; m++) order[cur[round[m]]++] = (uint32_t)m;
    free(cur);
    return R;
}

/* ------------------------------------------------------------------ */
/* synthetic generator -- host twin of the CUDA one      */
/* ------------------------------------------------------------------ */
typedef struct {
    uint64_t seed;
    uint64_t n_players;
    int32_t slots;        /* P: 6 for 3v3, 10 for 5v5 */
    int32_t mode;         /* rating column selector written into meta */
    uint32_t draw_thr;    /* P(draw) * 2^32 */
    uint32_t afk_thr;     /* P(afk gate) * 2^32 */
    int32_t zipf;         /* 0 uniform, 1 Zipf-Mandelbrot (r + q)^-0.8 */
    double zipf_a;        /* q^0.2 */
    double zipf_b;        /* (U + q)^0.2 - q^0.2 */
    double zipf_q;
} tso_gen;

static inline uint64_t mix64(uint64_t z)
{
    z += 0x9E3779B97F4A7C15ull;
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
}
static inline uint64_t rnd3(uint64_t seed, uint64_t a, uint64_t b, uint64_t c)
{
    return mix64(mix64(mix64(seed ^ a) + b * 0xD1342543DE82EF95ull) + c);
}
static inline int32_t latent_skill(uint64_t seed, uint64_t pid)
{
    uint64_t h = mix64(mix64(seed ^ 0x5EEDFACE5EEDFACEull) + pid);
    return (int32_t)((h & 0xFFFF) + ((h >> 16) & 0xFFFF) + ((h >> 32) & 0xFFFF) + (h >> 48)) - 2 * 65535;
}
static inline uint32_t sample_player(const tso_gen *g, uint64_t x)
{
    if (!g->zipf) return (uint32_t)(((unsigned __int128)x * g->n_players) >> 64);
    double u = (double)(x >> 11) * 0x1.0p-53;
    double y = u * g->zipf_b;
    y = y + g->zipf_a;
    double y2 = y * y;
    double y4 = y2 * y2;
    double y5 = y4 * y;
    double r = y5 - g->zipf_q;
    if (!(r >= 0.0)) r = 0.0;
    uint64_t k = (uint64_t)r;
    if (k >= g->n_players) k = g->n_players - 1;
    return (uint32_t)k;
}

void tso_gen_matches(const tso_gen *g, int64_t first, int64_t count, uint32_t *ids, uint32_t *meta)
{
    int P = g->slots, half = P / 2;
    for (int64_t i = 0; i < count; i++) {
        uint64_t m = (uint64_t)(first + i);
        uint32_t pid[TSO_MAXP];
        int64_t strength = 0;
        for (int s = 0; s < P; s++) {
            uint64_t draw = 0;
            for (;;) {
                uint32_t p = sample_player(g, rnd3(g->seed, m, (uint64_t)s, draw));
                int dup = 0;
                for (int k = 0; k < s; k++) dup |= (pid[k] == p);
                if (!dup) { pid[s] = p; break; }
                draw++;
            }
            int32_t sk = latent_skill(g->seed, pid[s]);
            strength += (s < half) ? sk : -sk;
            ids[i * P + s] = pid[s];
        }
        uint64_t x = rnd3(g->seed, m, 0xFFFFull, 0);
        int64_t thr = ((int64_t)1 << 31) + strength * 11585;
        int64_t lo = 85899346, hi = 4209067950ll; /* 0.02, 0.98 of 2^32 */
        if (thr < lo) thr = lo;
        if (thr > hi) thr = hi;
        uint32_t bits = ((int64_t)(x >> 32) < thr) ? META_WIN_A : META_WIN_B;
        uint64_t y = rnd3(g->seed, m, 0xFFFEull, 0);
        if ((uint32_t)y < g->draw_thr) bits = 0;
        if ((uint32_t)(y >> 32) < g->afk_thr) bits |= META_AFK;
        meta[i] = bits | ((uint32_t)g->mode << 4);
    }
}

/* initial single-pair table: mu = 1500 + 400 z, sigma = 500 * (2/3) */
void tso_gen_table(uint64_t seed, int64_t first, int64_t count, double *table)
{
    for (int64_t i = 0; i < count; i++) {
        uint64_t h = mix64(mix64(seed ^ 0x7AB1E7AB1E7AB1E5ull) + (uint64_t)(first + i));
        int32_t s4 = (int32_t)((h & 0xFFFF) + ((h >> 16) & 0xFFFF) + ((h >> 32) & 0xFFFF) + (h >> 48)) - 2 * 65535;
        double z = (double)s4 * (1.0 / 37837.0);
        double mu = 400.0 * z;
        mu = mu + 1500.0;
        table[i * 2] = mu;
        table[i * 2 + 1] = 500.0 * (2.0 / 3.0);
    }
}
