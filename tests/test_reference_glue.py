"""The compiled reference glue model (oracle/_ref/libglue_ref.so: 3-sim-1kb's unmodified glue_simulator.cpp behind
oracle/ref_glue.cc) loads and behaves like the class it is -- so that tests/test_gpu_glue.py, which holds the
engine's bd_glue_update to it, compares against something that works.  CPU only."""
import math

import numpy as np
import pytest

import bdb200 as bd
import checkers


@pytest.fixture(scope="module")
def cloud(oracle):
    rng = np.random.default_rng(2)
    period = (3.0, 3.2, 2.9)
    x = rng.uniform(0, 1, size=(600, 3)) * period
    dist = 0.3
    cand = oracle.nsearch_pairs(bd.BOX_PERIODIC, period, len(x), dist, x)
    return x, period, dist, cand


def make(cloud, max_glues, p_bind, p_unbind, seed):
    if not checkers.have_reference_glue():
        pytest.skip("oracle/_ref/libglue_ref.so not built (no /root/reference here)")
    x, period, dist, _ = cloud
    rate = lambda p: 1e300 if p >= 1 else -math.log1p(-p)  # noqa: E731
    return checkers.ReferenceGlue(max_glues, dist, rate(p_bind), rate(p_unbind), period, seed)


def pair_keys(p):
    p = np.sort(p.astype(np.uint64), axis=1)
    return np.sort((p[:, 0] << np.uint64(32)) | p[:, 1])


def test_reference_glue_binds_candidates_only_and_respects_the_capacity(cloud):
    x, period, dist, cand = cloud
    assert len(cand) > 200
    g = make(cloud, 50, 1.0, 0.0, seed=1)
    assert g.update(1.0, x) == 50 and g.update(1.0, x) == 50   # saturated, nothing unbinds
    got = pair_keys(g.pairs())
    assert len(np.unique(got)) == 50 and np.isin(got, pair_keys(cand)).all()
    g.close()
    # everything binds when there is room: the glued set IS the search result
    g = make(cloud, 10 * len(cand), 1.0, 0.0, seed=1)
    assert g.update(1.0, x) == len(cand)
    assert np.array_equal(pair_keys(g.pairs()), pair_keys(cand))
    # pairs pulled apart come off whatever the unbinding rate
    assert g.update(1.0, x + np.arange(len(x))[:, None] * 0.0) == len(cand)
    far = x.copy()
    far[::2] += 0.5 * np.array(period)
    kept = g.update(1.0, far)
    d = far[g.pairs()[:, 0]] - far[g.pairs()[:, 1]]
    d -= np.rint(d / np.array(period)) * np.array(period)
    assert kept == len(g.pairs()) and (np.sqrt((d * d).sum(axis=1)) <= dist).all()
    g.close()


def test_reference_glue_rates_and_early_return(cloud):
    x, period, dist, cand = cloud
    n = []
    for seed in range(200):
        g = make(cloud, 10 * len(cand), 0.1, 0.0, seed=seed)
        n.append(g.update(1.0, x))
        g.close()
    m = len(cand)
    assert abs(np.mean(n) - 0.1 * m) <= 5 * math.sqrt(0.1 * 0.9 * m / 200)
    g = make(cloud, 0, 1.0, 1.0, seed=3)
    assert g.update(1.0, x) == 0 and len(g.pairs()) == 0
    g.close()
