"""Kinetic glue update on the device (bd_glue_update) against the reference algorithm.  GPU only.

The reference (3-sim-1kb/src/simulation/glues/glue_simulator.cpp:35-80, reservoir_sampler.hpp:57-93) walks the
pairs of a host-side neighbour search through ONE sequential generator: unbond by distance or a Bernoulli draw,
bond free pairs by a Bernoulli draw, fill the free capacity with a uniform subset (reservoir sampling,
"Algorithm L").  The device version draws counter-based numbers per pair, so it is the same stochastic process
but not the same sample path.  What is held here:

* hard invariants, every update: pairs farther than max_distance are gone, no pair is glued twice, every new
  pair was a free pair within max_distance (the reference's pair set), the list never exceeds max_glues;
* rates: numbers unbound / bound per update are Binomial(glued, p_unbind) / Binomial(free, p_bind) -- mean
  within 5 standard errors over many independent updates of the same state;
* the saturated case: when more pairs pass than fit, each is chosen equally often (chi-square), which is what
  the reference's reservoir guarantees;
* against THE REFERENCE ITSELF -- the unmodified glue_simulator.cpp compiled into oracle/_ref/libglue_ref.so
  (oracle/ref_glue.cc, oracle/Makefile) and run with its own std::mt19937_64 on the same positions: the
  glue-count trajectory of a run of updates (mean at every update, spread at the stationary end), the pairs
  lost and gained per update, the stationary count, and the hit histogram of the saturated subset (the
  reference's reservoir passes the same chi-square test the engine is held to);
* max_glues = 0 leaves the list alone, as the reference's early return does (glue_simulator.cpp:37-39).
"""
import math

import numpy as np
import pytest

import bdb200 as bd
import scenarios as sc
from cases import sort_pairs

pytestmark = pytest.mark.gpu

DIST = 0.42


def keys(p):
    p = sort_pairs(p).astype(np.uint64)
    return (p[:, 0] << np.uint64(32)) | p[:, 1]


def box_distance(x, pairs, period):
    d = x[pairs[:, 0]] - x[pairs[:, 1]]
    L = np.asarray(period)
    d = d - np.rint(d * (1 / L)) * L
    return np.sqrt((d * d).sum(axis=1))


def reference_glue(scen, max_glues, p_bind, p_unbind, seed):
    """the reference class with rates such that -expm1(-rate * 1) = p (the driver's conversion, timestep 1)"""
    import checkers

    if not checkers.have_reference_glue():
        pytest.skip("oracle/_ref/libglue_ref.so not built (no /root/reference here)")

    def rate(p):
        return 1e300 if p >= 1 else -math.log1p(-p)

    return checkers.ReferenceGlue(max_glues, DIST, rate(p_bind), rate(p_unbind), scen.period, seed)


@pytest.fixture(scope="module")
def world(engine, oracle):
    scen = sc.chromatin_1kb(3, 400, seed=8)
    s = bd.System(engine, scen.n)
    ids = scen.apply(s)
    s.step(200, scen.timestep, scen.temperature, seed=2)  # a relaxed configuration
    x = s.get_positions()
    cand = sort_pairs(oracle.nsearch_pairs(bd.BOX_PERIODIC, scen.period, scen.n, DIST, x))
    glue_ff = ids["bond"][1]
    s.set_bond_min_image(glue_ff, True)
    return scen, s, glue_ff, x, cand


def test_invariants_and_rates(world):
    scen, s, ff, x, cand = world
    rng = np.random.default_rng(0)
    ck = keys(cand)
    start = cand[rng.choice(len(cand), size=len(cand) // 5, replace=False)]
    far = np.array([[0, 700], [5, 900]], dtype=np.uint32)  # two glued pairs that are too far apart
    assert (box_distance(x, far, scen.period) > DIST).all()
    start = np.concatenate([start, far])
    p_bind, p_unbind, cap = 0.03, 0.2, 10 * len(cand)
    free = len(cand) - (len(start) - len(far))
    m = len(start) - len(far)
    unbound, bound = [], []
    for u in range(300):
        s.set_bond_pairs(ff, start)
        a, b, n = s.glue_update(ff, DIST, p_bind, p_unbind, cap, seed=11, update_index=u)
        new = s.get_bond_pairs(ff)
        nk = keys(new)
        assert n == len(new) == len(start) - a + b and len(np.unique(nk)) == len(nk)
        assert np.isin(nk, ck).all()                      # only pairs of the reference's search set, so the far ones left
        assert (box_distance(x, new, scen.period) <= DIST).all()
        # bound pairs are free pairs: new ones, or ones that came apart in this very update (the reference binds
        # against the list AFTER unbond, glue_simulator.cpp:41-42,74)
        fresh = len(np.setdiff1d(nk, keys(start)))
        assert fresh <= b <= fresh + a
        unbound.append(a - len(far))
        bound.append(b)
    assert abs(np.mean(unbound) - m * p_unbind) <= 5 * math.sqrt(m * p_unbind * (1 - p_unbind) / 300)
    free_mean = free + m * p_unbind  # the pairs that just came apart are free again
    assert abs(np.mean(bound) - free_mean * p_bind) <= 5 * math.sqrt(free_mean * p_bind * (1 - p_bind) / 300)
    assert abs(np.var(unbound) / (m * p_unbind * (1 - p_unbind)) - 1) < 0.5  # binomial spread, not a constant


def test_saturated_binding_is_a_uniform_subset(world):
    scen, s, ff, x, cand = world
    ck = keys(cand)
    # p_bind = 1: every free pair passes; capacity 40 of them are chosen per update
    updates = 400
    hits = {"engine": np.zeros(len(cand)), "reference": np.zeros(len(cand))}
    for u in range(updates):
        s.set_bond_pairs(ff, np.zeros((0, 2), np.uint32))
        a, b, n = s.glue_update(ff, DIST, 1.0, 0.0, 40, seed=5, update_index=u)
        assert (a, b, n) == (0, 40, 40)
        hits["engine"][np.searchsorted(ck, keys(s.get_bond_pairs(ff)))] += 1
        ref = reference_glue(scen, 40, 1.0, 0.0, seed=900 + u)  # a fresh, empty simulator: one saturated update
        assert ref.update(1.0, x) == 40
        rk = keys(ref.pairs())
        assert np.isin(rk, ck).all()
        hits["reference"][np.searchsorted(ck, rk)] += 1
        ref.close()
    expect = updates * 40 / len(cand)
    dof = len(cand) - 1
    for who, h in hits.items():
        chi2 = ((h - expect) ** 2 / expect).sum()
        assert abs(chi2 - dof) <= 5 * math.sqrt(2 * dof), (who, chi2, dof)


def test_no_capacity_leaves_the_list_alone(world):
    """glue_simulator::update returns at once when max_glues == 0: nothing comes apart, not even pairs too far apart"""
    scen, s, ff, x, cand = world
    start = np.concatenate([cand[:50], np.array([[0, 700]], dtype=np.uint32)])
    s.set_bond_pairs(ff, start)
    assert s.glue_update(ff, DIST, 0.5, 0.5, 0, seed=3, update_index=0) == (0, 0, len(start))
    assert np.array_equal(s.get_bond_pairs(ff), start)
    ref = reference_glue(scen, 0, 0.5, 0.5, seed=1)
    assert ref.update(1.0, x) == 0
    ref.close()


@pytest.mark.parametrize("regime", ["capacity-limited", "rate-limited"])
def test_glue_trajectory_matches_the_reference_class(world, regime):
    """capacity-limited: more pairs pass the binding draw than fit, the list sits at max_glues (the reservoir decides
    who gets in); rate-limited: max_glues is never reached, the count fluctuates around the balance of the two rates"""
    scen, s, ff, x, cand = world
    p_bind, p_unbind = 0.02, 0.1
    cap = len(cand) // 8 if regime == "capacity-limited" else 10 * len(cand)
    runs, length = 24, 40
    count = {"engine": np.zeros((runs, length)), "reference": np.zeros((runs, length))}
    lost = {"engine": np.zeros((runs, length)), "reference": np.zeros((runs, length))}
    gained = {"engine": np.zeros((runs, length)), "reference": np.zeros((runs, length))}
    ck = keys(cand)

    def record(who, r, u, before, after):
        count[who][r, u] = len(after)
        lost[who][r, u] = len(np.setdiff1d(before, after))
        gained[who][r, u] = len(np.setdiff1d(after, before))

    for r in range(runs):
        s.set_bond_pairs(ff, np.zeros((0, 2), np.uint32))
        ref = reference_glue(scen, cap, p_bind, p_unbind, seed=100 + r)
        before_e = before_r = np.zeros(0, dtype=np.uint64)
        for u in range(length):
            n = s.glue_update(ff, DIST, p_bind, p_unbind, cap, seed=1000 + r, update_index=u)[2]
            after_e = keys(s.get_bond_pairs(ff))
            assert n == len(after_e)
            assert ref.update(1.0, x) == len(ref.pairs())
            after_r = keys(ref.pairs())
            assert np.isin(after_r, ck).all() and len(after_r) <= cap  # the reference obeys the engine's invariants too
            record("engine", r, u, before_e, after_e)
            record("reference", r, u, before_r, after_r)
            before_e, before_r = after_e, after_r
        ref.close()
    # the same relaxation towards the same stationary state: means over runs at every update, for the count and
    # for the pairs lost / gained in that update (set differences, measured the same way on both sides)
    for what in (count, lost, gained):
        got, want = what["engine"], what["reference"]
        se = np.sqrt(got.var(axis=0) / runs + want.var(axis=0) / runs) + 0.5
        assert (np.abs(got.mean(axis=0) - want.mean(axis=0)) <= 5 * se).all()
    got, want = count["engine"], count["reference"]
    assert got[:, -1].mean() > 10  # the list did fill up
    # stationary end: mean within 5 standard errors
    tail_e, tail_r = got[:, -5:], want[:, -5:]
    assert abs(tail_e.mean() - tail_r.mean()) <= 5 * math.sqrt(tail_e.var() / runs + tail_r.var() / runs) + 0.5
    if regime == "capacity-limited":
        # both sit at max_glues nearly always (now and then fewer pairs pass the draw than came apart)
        assert (tail_r == cap).mean() >= 0.9 and (tail_e == cap).mean() >= 0.9
        assert tail_e.max() <= cap and tail_r.max() <= cap
    else:
        # binomial fluctuation around the rate balance: the same spread on both sides, and never near the capacity
        assert tail_r.std() > 3 and abs(tail_e.std() / tail_r.std() - 1) < 0.6
        expect = len(cand) * p_bind / (p_bind + p_unbind - p_bind * p_unbind)  # balance of bind and unbind per update
        assert abs(tail_r.mean() / expect - 1) < 0.1 and abs(tail_e.mean() / expect - 1) < 0.1
