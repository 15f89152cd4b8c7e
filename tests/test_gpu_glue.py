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
* the glue-count trajectory of a run of updates has the mean and spread of the reference algorithm restated
  in numpy (`reference_update`, run with its own generator on the same positions).
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


def reference_update(glued, candidates, dist_of, p_bind, p_unbind, max_glues, rng):
    """glue_simulator::update restated: `glued` a set of (i, j), `candidates` the search result in emission
    order, dist_of(pair) the box distance.  Reservoir sampling is Algorithm L as in reservoir_sampler.hpp."""
    glued = {g for g in glued if not (dist_of(g) > DIST or rng.random() < p_unbind)}
    capacity = max_glues - len(glued)
    samples, population, skip, acceptance = [], 0, 0, 1.0

    def update_skip():
        nonlocal skip, acceptance
        acceptance *= rng.random() ** (1.0 / capacity)
        skip = rng.geometric(acceptance) - 1 if acceptance < 1 else 0

    for pair in candidates:
        if pair in glued or not (rng.random() < p_bind) or capacity == 0:
            continue
        population += 1
        if population <= capacity:
            samples.append(pair)
            continue
        if population == capacity + 1:
            update_skip()
        if skip:
            skip -= 1
            continue
        samples[rng.integers(capacity)] = pair
        update_skip()
    return glued | set(samples)


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
    sub = cand[:240]
    # p_bind = 1: every free pair passes; capacity 40 of them are chosen per update
    hits = np.zeros(len(cand))
    updates = 400
    for u in range(updates):
        s.set_bond_pairs(ff, np.zeros((0, 2), np.uint32))
        a, b, n = s.glue_update(ff, DIST, 1.0, 0.0, 40, seed=5, update_index=u)
        assert (a, b, n) == (0, 40, 40)
        hits[np.searchsorted(keys(cand), keys(s.get_bond_pairs(ff)))] += 1
    expect = updates * 40 / len(cand)
    chi2 = ((hits - expect) ** 2 / expect).sum()
    dof = len(cand) - 1
    assert abs(chi2 - dof) <= 5 * math.sqrt(2 * dof), (chi2, dof)
    del sub


def test_glue_count_trajectory_matches_the_reference_algorithm(world):
    scen, s, ff, x, cand = world
    p_bind, p_unbind, cap = 0.02, 0.1, len(cand) // 8
    dist = dict(zip(map(tuple, cand.tolist()), box_distance(x, cand, scen.period)))
    runs, length = 24, 40
    got, want = np.zeros((runs, length)), np.zeros((runs, length))
    order = [tuple(p) for p in cand.tolist()]
    for r in range(runs):
        s.set_bond_pairs(ff, np.zeros((0, 2), np.uint32))
        glued = set()
        rng = np.random.default_rng(100 + r)
        for u in range(length):
            got[r, u] = s.glue_update(ff, DIST, p_bind, p_unbind, cap, seed=1000 + r, update_index=u)[2]
            glued = reference_update(glued, order, lambda g: dist[g], p_bind, p_unbind, cap, rng)
            want[r, u] = len(glued)
    # same relaxation towards the same stationary count; compare means over runs at every update
    se = np.sqrt(got.var(axis=0) / runs + want.var(axis=0) / runs) + 0.5
    assert (np.abs(got.mean(axis=0) - want.mean(axis=0)) <= 5 * se).all()
    assert got[:, -1].mean() > 5 * got[:, 0].mean() * 0 + 10  # the list did fill up
    assert abs(got[:, -5:].std() / max(want[:, -5:].std(), 1e-9) - 1) < 0.6
