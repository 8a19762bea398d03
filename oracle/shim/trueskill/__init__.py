"""ORACLE O1 -- test infrastructure, NOT product code.

A from-scratch stand-in for the third-party ``trueskill==0.4.4`` package
(pinned at /root/reference requirements.txt:6, backend ``mpmath==0.19`` pinned
at requirements.txt:2).  That package is not in /root/reference, is not
installed in this image and cannot be fetched (no network), so its published
algorithm -- the TrueSkill factor graph of Herbrich, Minka & Graepel (NIPS
2006) with the message schedule the package uses for two or more teams -- is
restated here behind the same public names, so that the UNMODIFIED reference
``rater.py`` imports and runs against it:

    rater.py:6      import trueskill
    rater.py:30-37  trueskill.TrueSkill(backend="mpmath", mu, sigma, beta, tau, draw_probability)
    rater.py:121    env.create_rating(mu, sigma)
    rater.py:141    env.quality(matchup)
    rater.py:144    env.rate(matchup_shared, ranks=[...])
    rater.py:161    env.rate(matchup, ranks=[...])

Parity status: the reference's own tests (worker_test.py:136-142,165) hold no
numeric golden value for rate()/quality(): PARITY UNPINNED by the reference.
This shim is pinned instead on the public TrueSkill known-answer vectors
(tests/test_oracle_kat.py).

Only tests/, __graft_entry__.smoke() and bench.py's CPU-baseline legs may
import this module.  The product path (analyzer_b200) never does.
"""
from __future__ import annotations

import math

__all__ = ["TrueSkill", "Rating", "rate", "quality", "setup", "global_env",
           "MU", "SIGMA", "BETA", "TAU", "DRAW_PROBABILITY", "DELTA"]

MU = 25.0
SIGMA = MU / 3
BETA = SIGMA / 2
TAU = SIGMA / 100
DRAW_PROBABILITY = 0.10
DELTA = 0.0001

_INF = float("inf")


# --------------------------------------------------------------------------
# Normal-distribution primitives per backend
# --------------------------------------------------------------------------
def _make_erfcinv(erfc, m):
    """Inverse complementary error function the way the package computes it:
    a rational first guess followed by two Newton corrections on ``erfc``."""
    def erfcinv(y):
        if y >= 2:
            return -100.0
        if y <= 0:
            return 100.0
        lower = y < 1
        if not lower:
            y = 2 - y
        t = m.sqrt(-2 * m.log(y / 2.0))
        x = -0.70711 * ((2.30753 + t * 0.27061) / (1.0 + t * (0.99229 + t * 0.04481)) - t)
        for _ in range(2):
            err = erfc(x) - y
            x += err / (1.12837916709551257 * m.exp(-(x ** 2)) - x * err)
        return x if lower else -x
    return erfcinv


def _make_ppf(erfc, m):
    erfcinv = _make_erfcinv(erfc, m)

    def ppf(x, mu=0, sigma=1):
        return mu - sigma * m.sqrt(2) * erfcinv(2 * x)
    return ppf


def _backend(name):
    """Return (cdf, pdf, ppf) for a backend name (None = plain ``math``)."""
    if name is None:
        def cdf(x, mu=0, sigma=1):
            return 0.5 * math.erfc(-(x - mu) / (sigma * math.sqrt(2)))

        def pdf(x, mu=0, sigma=1):
            return (1 / math.sqrt(2 * math.pi) * abs(sigma)) * math.exp(-(((x - mu) / abs(sigma)) ** 2 / 2))
        return cdf, pdf, _make_ppf(math.erfc, math)
    if name == "mpmath":
        import mpmath
        return mpmath.ncdf, mpmath.npdf, _make_ppf(mpmath.erfc, mpmath)
    if name == "scipy":
        from scipy.stats import norm
        return norm.cdf, norm.pdf, norm.ppf
    raise ValueError("%r backend is not defined" % (name,))


# --------------------------------------------------------------------------
# Gaussians in natural (precision) parameters
# --------------------------------------------------------------------------
class Gaussian(object):
    """N(mu, sigma^2) stored as pi = sigma^-2 and tau = pi*mu."""

    __slots__ = ("pi", "tau")

    def __init__(self, mu=None, sigma=None, pi=0, tau=0):
        if mu is not None:
            if sigma is None:
                raise TypeError("sigma argument is needed")
            if sigma == 0:
                raise ValueError("sigma**2 should be greater than 0")
            pi = sigma ** -2
            tau = pi * mu
        self.pi = pi
        self.tau = tau

    @property
    def mu(self):
        return self.pi and self.tau / self.pi

    @property
    def sigma(self):
        return math.sqrt(1 / self.pi) if self.pi else _INF

    def __mul__(self, other):
        return Gaussian(pi=self.pi + other.pi, tau=self.tau + other.tau)

    def __truediv__(self, other):
        return Gaussian(pi=self.pi - other.pi, tau=self.tau - other.tau)

    __div__ = __truediv__

    def __repr__(self):
        return "N(mu=%.3f, sigma=%.3f)" % (self.mu, self.sigma)


class Rating(Gaussian):
    """A player's skill belief.  Behaves like a (mu, sigma) pair."""

    def __init__(self, mu=None, sigma=None):
        if isinstance(mu, tuple):
            mu, sigma = mu
        elif isinstance(mu, Gaussian):
            mu, sigma = mu.mu, mu.sigma
        if mu is None:
            mu = global_env().mu
        if sigma is None:
            sigma = global_env().sigma
        super(Rating, self).__init__(mu, sigma)

    def __int__(self):
        return int(self.mu)

    def __float__(self):
        return float(self.mu)

    def __iter__(self):
        return iter((self.mu, self.sigma))

    def __repr__(self):
        return "trueskill.Rating(mu=%.3f, sigma=%.3f)" % (self.mu, self.sigma)


# --------------------------------------------------------------------------
# Factor graph: variables hold a marginal and the last message of each factor
# --------------------------------------------------------------------------
class _Variable(Gaussian):
    def __init__(self):
        super(_Variable, self).__init__(pi=0, tau=0)
        self.messages = {}

    def _set(self, other):
        change = max(abs(self.tau - other.tau), math.sqrt(abs(self.pi - other.pi)))
        self.pi, self.tau = other.pi, other.tau
        return change

    def put_message(self, factor, pi, tau):
        msg = Gaussian(pi=pi, tau=tau)
        old = self.messages[factor]
        self.messages[factor] = msg
        return self._set(self / old * msg)

    def put_value(self, factor, pi=0, tau=0, value=None):
        if value is None:
            value = Gaussian(pi=pi, tau=tau)
        old = self.messages[factor]
        self.messages[factor] = value * old / self
        return self._set(value)


class _Factor(object):
    def __init__(self, variables):
        self.vars = variables
        for v in variables:
            v.messages[self] = Gaussian()


class _Prior(_Factor):
    """skill ~ N(mu, sigma^2 + tau_dyn^2)."""

    def __init__(self, var, rating, dynamic):
        super(_Prior, self).__init__([var])
        self.var, self.rating, self.dynamic = var, rating, dynamic

    def down(self):
        s = math.sqrt(self.rating.sigma ** 2 + self.dynamic ** 2)
        return self.var.put_value(self, value=Gaussian(self.rating.mu, s))


class _Likelihood(_Factor):
    """performance ~ N(skill, beta^2)."""

    def __init__(self, skill, perf, variance):
        super(_Likelihood, self).__init__([skill, perf])
        self.skill, self.perf, self.variance = skill, perf, variance

    def _shrink(self, g):
        return 1.0 / (1.0 + self.variance * g.pi)

    def down(self):
        msg = self.skill / self.skill.messages[self]
        a = self._shrink(msg)
        return self.perf.put_message(self, a * msg.pi, a * msg.tau)

    def up(self):
        msg = self.perf / self.perf.messages[self]
        a = self._shrink(msg)
        return self.skill.put_message(self, a * msg.pi, a * msg.tau)


class _Sum(_Factor):
    """total = sum(coeff_i * term_i)."""

    def __init__(self, total, terms, coeffs):
        super(_Sum, self).__init__([total] + list(terms))
        self.total, self.terms, self.coeffs = total, list(terms), list(coeffs)

    def _send(self, target, inputs, coeffs):
        var_sum = 0
        mean = 0
        for v, c in zip(inputs, coeffs):
            rest = v / v.messages[self]
            mean += c * rest.mu
            if var_sum == _INF:
                continue
            try:
                var_sum += c ** 2 / rest.pi
            except ZeroDivisionError:
                var_sum = _INF
        pi = 1.0 / var_sum
        return target.put_message(self, pi, pi * mean)

    def down(self):
        return self._send(self.total, self.terms, self.coeffs)

    def up(self, index=0):
        c0 = self.coeffs[index]
        coeffs = []
        for i, c in enumerate(self.coeffs):
            try:
                if i == index:
                    coeffs.append(1.0 / c0)
                else:
                    coeffs.append(-c / c0)
            except ZeroDivisionError:
                coeffs.append(0.0)
        inputs = list(self.terms)
        inputs[index] = self.total
        return self._send(self.terms[index], inputs, coeffs)


class _Truncate(_Factor):
    """Moment-matched truncation of the team-performance difference."""

    def __init__(self, var, v_func, w_func, margin):
        super(_Truncate, self).__init__([var])
        self.var, self.v_func, self.w_func, self.margin = var, v_func, w_func, margin

    def up(self):
        rest = self.var / self.var.messages[self]
        root = math.sqrt(rest.pi)
        t, eps = rest.tau / root, self.margin * root
        v = self.v_func(t, eps)
        w = self.w_func(t, eps)
        denom = 1.0 - w
        return self.var.put_value(self, rest.pi / denom, (rest.tau + root * v) / denom)


def calc_draw_probability(draw_margin, size, env=None):
    env = env or global_env()
    return 2 * env.cdf(draw_margin / (math.sqrt(size) * env.beta)) - 1


def calc_draw_margin(draw_probability, size, env=None):
    env = env or global_env()
    return env.ppf((draw_probability + 1) / 2.0) * math.sqrt(size) * env.beta


class TrueSkill(object):
    """A TrueSkill environment (rater.py:30-37 constructs one)."""

    def __init__(self, mu=MU, sigma=SIGMA, beta=BETA, tau=TAU,
                 draw_probability=DRAW_PROBABILITY, backend=None):
        self.mu, self.sigma, self.beta, self.tau = mu, sigma, beta, tau
        self.draw_probability = draw_probability
        self.backend = backend
        self.cdf, self.pdf, self.ppf = _backend(backend)

    # -- public helpers ----------------------------------------------------
    def create_rating(self, mu=None, sigma=None):
        return Rating(self.mu if mu is None else mu, self.sigma if sigma is None else sigma)

    Rating = create_rating

    def expose(self, rating):
        return rating.mu - (self.mu / self.sigma) * rating.sigma

    def make_as_global(self):
        return setup(env=self)

    # -- truncated-Gaussian corrections (Appendix A) -------------------------
    def _fp_error(self):
        if self.backend is None:
            hint = "Cannot calculate correctly, set backend to 'mpmath'"
        else:
            hint = "%r backend cannot calculate correctly" % (self.backend,)
        return FloatingPointError(hint)

    def v_win(self, t, eps):
        x = t - eps
        mass = self.cdf(x)
        return (self.pdf(x) / mass) if mass else -x

    def w_win(self, t, eps):
        x = t - eps
        v = self.v_win(t, eps)
        w = v * (v + x)
        if 0 < w < 1:
            return w
        raise self._fp_error()

    def v_draw(self, t, eps):
        at = abs(t)
        a, b = eps - at, -eps - at
        mass = self.cdf(a) - self.cdf(b)
        top = self.pdf(b) - self.pdf(a)
        return ((top / mass) if mass else a) * (-1 if t < 0 else +1)

    def w_draw(self, t, eps):
        at = abs(t)
        a, b = eps - at, -eps - at
        mass = self.cdf(a) - self.cdf(b)
        if not mass:
            raise self._fp_error()
        v = self.v_draw(at, eps)
        return (v ** 2) + (a * self.pdf(a) - b * self.pdf(b)) / mass

    # -- argument checking ---------------------------------------------------
    @staticmethod
    def _check_groups(groups):
        if len(groups) < 2:
            raise ValueError("Need multiple rating groups")
        if not all(groups):
            raise ValueError("Each group must contain multiple ratings")
        kinds = set(isinstance(g, dict) for g in groups)
        if len(kinds) != 1:
            raise TypeError("All groups should be same type")
        if kinds.pop():
            keys = [list(g.keys()) for g in groups]
            groups = [tuple(g.values()) for g in groups]
            return groups, keys
        return [tuple(g) for g in groups], None

    @staticmethod
    def _check_weights(weights, groups, keys):
        if weights is None:
            return [(1,) * len(g) for g in groups]
        if isinstance(weights, dict):
            table = [[1] * len(g) for g in groups]
            for (gi, key), w in weights.items():
                pos = keys[gi].index(key) if keys is not None else key
                table[gi][pos] = w
            return [tuple(r) for r in table]
        return [tuple(w) for w in weights]

    # -- the update -----------------------------------------------------------
    def rate(self, rating_groups, ranks=None, weights=None, min_delta=DELTA):
        groups, keys = self._check_groups(rating_groups)
        weights = self._check_weights(weights, groups, keys)
        n_groups = len(groups)
        if ranks is None:
            ranks = list(range(n_groups))
        elif len(ranks) != n_groups:
            raise ValueError("Wrong ranks")
        # stable sort of the teams by rank: lower rank (the winner) first
        perm = sorted(range(n_groups), key=lambda i: ranks[i])
        s_groups = [groups[i] for i in perm]
        s_ranks = [ranks[i] for i in perm]
        s_weights = [tuple(max(min_delta, w) for w in weights[i]) for i in perm]

        flat_ratings = [r for g in s_groups for r in g]
        flat_weights = [w for g in s_weights for w in g]
        sizes = []
        acc = 0
        for g in s_groups:
            acc += len(g)
            sizes.append(acc)

        skill = [_Variable() for _ in flat_ratings]
        perf = [_Variable() for _ in flat_ratings]
        team = [_Variable() for _ in s_groups]
        diff = [_Variable() for _ in s_groups[:-1]]

        priors = [_Prior(v, r, self.tau) for v, r in zip(skill, flat_ratings)]
        likes = [_Likelihood(s, p, self.beta ** 2) for s, p in zip(skill, perf)]
        team_sums = []
        for ti, tv in enumerate(team):
            lo = sizes[ti - 1] if ti > 0 else 0
            hi = sizes[ti]
            team_sums.append(_Sum(tv, perf[lo:hi], flat_weights[lo:hi]))
        diff_sums = [_Sum(dv, team[i:i + 2], [+1, -1]) for i, dv in enumerate(diff)]
        truncs = []
        for i, dv in enumerate(diff):
            dp = self.draw_probability
            if callable(dp):
                dp = dp(s_groups[i][0], s_groups[i + 1][0], self)
            size = len(s_groups[i]) + len(s_groups[i + 1])
            margin = calc_draw_margin(dp, size, self)
            if s_ranks[i] == s_ranks[i + 1]:
                fv, fw = self.v_draw, self.w_draw
            else:
                fv, fw = self.v_win, self.w_win
            truncs.append(_Truncate(dv, fv, fw, margin))

        # downward sweep
        for f in priors:
            f.down()
        for f in likes:
            f.down()
        for f in team_sums:
            f.down()
        # inner loop over the difference chain until messages settle
        n_diff = len(diff_sums)
        for _ in range(10):
            if n_diff == 1:
                diff_sums[0].down()
                delta = truncs[0].up()
            else:
                delta = 0
                for i in range(n_diff - 1):
                    diff_sums[i].down()
                    delta = max(delta, truncs[i].up())
                    diff_sums[i].up(1)
                for i in range(n_diff - 1, 0, -1):
                    diff_sums[i].down()
                    delta = max(delta, truncs[i].up())
                    diff_sums[i].up(0)
            if delta <= min_delta:
                break
        # upward sweep
        diff_sums[0].up(0)
        diff_sums[n_diff - 1].up(1)
        for f in team_sums:
            for i in range(len(f.terms)):
                f.up(i)
        for f in likes:
            f.up()

        out_sorted = []
        for ti in range(n_groups):
            lo = sizes[ti - 1] if ti > 0 else 0
            hi = sizes[ti]
            out_sorted.append(tuple(Rating(float(v.mu), float(v.sigma)) for v in skill[lo:hi]))
        out = [None] * n_groups
        for pos, i in enumerate(perm):
            out[i] = out_sorted[pos]
        if keys is None:
            return out
        return [dict(zip(keys[i], g)) for i, g in enumerate(out)]

    def quality(self, rating_groups, weights=None):
        """Draw probability of the match-up (closed form of the package's
        matrix expression exp(-m'A (B)^-1 A'm / 2) * sqrt(det(b2 A'A)/det(B)),
        B = b2 A'A + A'SA, evaluated in plain floats like the package does)."""
        groups, keys = self._check_groups(rating_groups)
        weights = self._check_weights(weights, groups, keys)
        mus = [float(r.mu) for g in groups for r in g]
        var = [float(r.sigma) ** 2 for g in groups for r in g]
        wts = [float(w) for g in weights for w in g]
        n_groups = len(groups)
        n = len(mus)
        # comparison matrix A^T: one row per adjacent team pair
        rows = []
        start = 0
        for gi in range(n_groups - 1):
            row = [0.0] * n
            a_len = len(groups[gi])
            b_len = len(groups[gi + 1])
            for k in range(a_len):
                row[start + k] = wts[start + k]
            for k in range(b_len):
                row[start + a_len + k] = -wts[start + a_len + k]
            rows.append(row)
            start += a_len
        m = len(rows)
        b2 = self.beta ** 2
        ata = [[b2 * sum(rows[i][k] * rows[j][k] for k in range(n)) for j in range(m)] for i in range(m)]
        atsa = [[sum(rows[i][k] * var[k] * rows[j][k] for k in range(n)) for j in range(m)] for i in range(m)]
        mid = [[ata[i][j] + atsa[i][j] for j in range(m)] for i in range(m)]
        end = [sum(rows[i][k] * mus[k] for k in range(n)) for i in range(m)]
        sol = _solve(mid, end)
        e_arg = -0.5 * sum(end[i] * sol[i] for i in range(m))
        s_arg = _det(ata) / _det(mid)
        return math.exp(e_arg) * math.sqrt(s_arg)


def _det(a):
    a = [row[:] for row in a]
    n = len(a)
    d = 1.0
    for c in range(n):
        p = max(range(c, n), key=lambda r: abs(a[r][c]))
        if a[p][c] == 0:
            return 0.0
        if p != c:
            a[c], a[p] = a[p], a[c]
            d = -d
        d *= a[c][c]
        for r in range(c + 1, n):
            f = a[r][c] / a[c][c]
            for k in range(c, n):
                a[r][k] -= f * a[c][k]
    return d


def _solve(a, b):
    n = len(a)
    a = [row[:] + [b[i]] for i, row in enumerate(a)]
    for c in range(n):
        p = max(range(c, n), key=lambda r: abs(a[r][c]))
        a[c], a[p] = a[p], a[c]
        for r in range(c + 1, n):
            f = a[r][c] / a[c][c]
            for k in range(c, n + 1):
                a[r][k] -= f * a[c][k]
    x = [0.0] * n
    for r in range(n - 1, -1, -1):
        x[r] = (a[r][n] - sum(a[r][k] * x[k] for k in range(r + 1, n))) / a[r][r]
    return x


# --------------------------------------------------------------------------
# module-level convenience API
# --------------------------------------------------------------------------
_global = []


def global_env():
    if not _global:
        setup()
    return _global[0]


def setup(mu=MU, sigma=SIGMA, beta=BETA, tau=TAU, draw_probability=DRAW_PROBABILITY,
          backend=None, env=None):
    if env is None:
        env = TrueSkill(mu, sigma, beta, tau, draw_probability, backend)
    del _global[:]
    _global.append(env)
    return env


def rate(rating_groups, ranks=None, weights=None, min_delta=DELTA):
    return global_env().rate(rating_groups, ranks, weights, min_delta)


def quality(rating_groups, weights=None):
    return global_env().quality(rating_groups, weights)


def rate_1vs1(rating1, rating2, drawn=False, min_delta=DELTA, env=None):
    env = env or global_env()
    teams = env.rate([(rating1,), (rating2,)], [0, 0 if drawn else 1], min_delta=min_delta)
    return teams[0][0], teams[1][0]


def quality_1vs1(rating1, rating2, env=None):
    env = env or global_env()
    return env.quality([(rating1,), (rating2,)])
