"""Pins oracle O1 (oracle/shim/trueskill) on the public TrueSkill known-answer
vectors (published to 3 decimals by the trueskill
package's own test-suite) and on the loose asserts of the reference's tests
(worker_test.py:67-189), which the shim + port must satisfy unchanged."""
import math
import os
import sys

import pytest

from oracle import rater_port as rp   # puts the shim on sys.path
import trueskill as ts

TOL = 1e-3   # the published vectors carry 3 decimals; the package's own tests compare with |a-b| < 1e-3


def _close(rating, mu, sigma):
    assert abs(rating.mu - mu) < TOL and abs(rating.sigma - sigma) < TOL, (rating, mu, sigma)


@pytest.fixture(params=[None, "mpmath"])
def env(request):
    return ts.TrueSkill(backend=request.param)


def test_1v1(env):
    r = env.create_rating
    a, b = env.rate([(r(),), (r(),)])
    _close(a[0], 29.396, 7.171)
    _close(b[0], 20.604, 7.171)
    a, b = env.rate([(r(),), (r(),)], ranks=[0, 0])
    _close(a[0], 25.000, 6.458)
    _close(b[0], 25.000, 6.458)
    a, b = env.rate([(r(25, 25 / 3.),), (r(50, 12.5),)], ranks=[0, 0])
    _close(a[0], 31.662, 7.137)
    _close(b[0], 35.010, 7.910)


def test_teams(env):
    r = env.create_rating
    a, b = env.rate([(r(), r()), (r(), r())])
    _close(a[0], 28.108, 7.774)
    _close(b[1], 21.892, 7.774)
    a, b = env.rate([(r(),) * 4, (r(),) * 4])
    _close(a[0], 27.198, 8.059)
    _close(b[0], 22.802, 8.059)
    a, b = env.rate([(r(28, 7), r(27, 6), r(26, 5)), (r(30, 4), r(31, 3))])
    for got, want in zip(a + b, [(28.658, 6.770), (27.484, 5.856), (26.336, 4.917),
                                 (29.785, 3.958), (30.879, 2.983)]):
        _close(got, *want)
    a, b = env.rate([(r(),), (r(), r())])
    _close(a[0], 33.730, 7.317)
    _close(b[0], 16.270, 7.317)
    a, b = env.rate([(r(),), (r(), r())], ranks=[0, 0])
    _close(a[0], 31.660, 7.138)
    _close(b[1], 18.340, 7.138)


def test_rank_sorting_and_three_teams(env):
    r = env.create_rating
    # loser listed first: the result comes back in input order
    a, b = env.rate([(r(),), (r(),)], ranks=[1, 0])
    _close(a[0], 20.604, 7.171)
    _close(b[0], 29.396, 7.171)
    a, b, c = env.rate([(r(),), (r(),), (r(),)])
    _close(a[0], 31.675, 6.656)
    _close(b[0], 25.000, 6.208)
    _close(c[0], 18.325, 6.656)


def test_quality(env):
    r = env.create_rating
    assert abs(env.quality([(r(),), (r(),)]) - 0.447) < TOL
    assert abs(env.quality([(r(28, 7), r(27, 6), r(26, 5)), (r(30, 4), r(31, 3))]) - 0.254) < TOL
    assert abs(env.quality([(r(25, 25 / 3.),), (r(50, 12.5),)]) - 0.110) < TOL


def test_errors(env):
    r = env.create_rating
    with pytest.raises(ValueError):
        env.rate([(r(),)])
    with pytest.raises(ValueError):
        env.rate([(r(),), ()])
    with pytest.raises(ValueError):
        env.create_rating(25, 0)


def test_float_backend_raises_where_mpmath_does_not():
    """rater.py:8 -- the mpmath backend exists because the float one overflows."""
    flt = ts.TrueSkill(backend=None, mu=1500, sigma=1000, beta=1000.0, tau=10.0, draw_probability=0)
    big, small = flt.create_rating(400000, 10), flt.create_rating(0, 10)
    with pytest.raises(FloatingPointError):
        flt.rate([(small,), (big,)])
    mp = rp.make_env()
    out = mp.rate([(mp.create_rating(400000, 10),), (mp.create_rating(0, 10),)], ranks=[1, 0])
    assert out[0][0].mu < 400000


def test_analyzer_env_vectors():
    """Derived (not reference-published) analyzer-env vectors, SURVEY Appendix B."""
    env = rp.make_env()
    seed = (1979.5454545454545, 500)
    w, l = env.rate([(env.create_rating(*seed),) * 3, (env.create_rating(*seed),) * 3], ranks=[0, 1])
    assert abs(w[0].mu - 2052.4082367567416) < 1e-9 and abs(w[0].sigma - 494.7635950314357) < 1e-9
    assert abs(l[0].mu - 1906.6826723341674) < 1e-9
    assert abs(env.quality([(env.create_rating(*seed),) * 3] * 2) - math.sqrt(0.8)) < 1e-12
    w, l = env.rate([(env.create_rating(2000, 100),) * 3] * 2, ranks=[0, 1])
    assert abs(w[0].mu - 2003.273434268004) < 1e-9 and abs(w[0].sigma - 100.44543109615816) < 1e-9
    d, _ = env.rate([(env.create_rating(2000, 100),) * 3] * 2, ranks=[1, 1])
    assert abs(d[0].mu - 2000.0) < 1e-9 and abs(d[0].sigma - 100.41498062126155) < 1e-9


def test_vst_points_and_seed():
    """worker_test.py:67-113 (exact pin of the seed policy)."""
    assert sorted(rp.vst_points) == list(range(-1, 30))
    mu, sigma = rp.seed_from_fields(None, None, 15)
    assert 1300 < mu - sigma < 1700
    for ranked, blitz in [(2500, None), (2500, 100), (100, 2500), (None, 2500), (2500, 0), (0, 2500)]:
        mu, sigma = rp.seed_from_fields(ranked, blitz, 0)
        assert mu - sigma == 2500
    with pytest.raises(KeyError):
        rp.seed_from_fields(None, None, 30)
    with pytest.raises(KeyError):
        rp.seed_from_fields(0, None, None)


def _reference_style_match(went_afk, shared):
    g_player = lambda: rp.new_player(skill_tier=15, trueskill_mu=shared and 2000,
                                     trueskill_sigma=shared and 100)
    part = lambda: rp.new_participant(g_player(), went_afk=went_afk)
    return rp.new_match("ranked", [(True, [part()] * 3), (False, [part()] * 3)])


def test_port_satisfies_reference_test_asserts():
    """worker_test.py:115-189 restated against the port (same loose asserts)."""
    m = _reference_style_match(False, None)
    rp.rate_match(m)
    p0, p1 = m.rosters[0].participants[0].player[0], m.rosters[1].participants[0].player[0]
    assert p0.trueskill_mu is not None and p0.trueskill_ranked_mu is not None
    assert p0.trueskill_ranked_sigma < p0.trueskill_ranked_mu
    assert 500 < p0.trueskill_ranked_mu < 2500
    assert p0.trueskill_casual_mu is None
    assert p0.trueskill_mu > p1.trueskill_mu and p0.trueskill_ranked_mu > p1.trueskill_ranked_mu
    m = _reference_style_match(False, True)
    rp.rate_match(m)
    assert 1800 < m.rosters[0].participants[0].player[0].trueskill_ranked_mu < 2200
    m = _reference_style_match(True, None)
    rp.rate_match(m)
    assert m.rosters[0].participants[0].player[0].trueskill_mu is None
    assert m.rosters[0].participants[0].participant_items[0].any_afk is True
    assert m.trueskill_quality == 0
