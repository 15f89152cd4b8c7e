"""The production noise of the step kernel (Philox4x32-10 + single-precision Box-Muller, bd_device.cuh:
philox_normals) against the exact standard normal, on 3e8 draws made on the device exactly as the kernel
makes them (bd_noise_statistics).  GPU only.

The reference draws double-precision ziggurat normals (MD/basic_types/ziggurat.hpp:262-322); the two streams
differ value by value, so the generator is pinned by its distribution.  Tolerances are 5 standard errors of
each estimator (all moments of N(0,1) are known in closed form: E z^2k = (2k-1)!!), tails by their exact
Gaussian mass with a Poisson band.
"""
import math

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

BEADS, STEPS = 1_000_000, 100  # 3e8 normals


def gauss_tail(k):
    return math.erfc(k / math.sqrt(2.0))  # P(|z| > k)


@pytest.fixture(scope="module")
def stats(engine):
    return engine.noise_statistics(seed=20221, first_step=1, steps=STEPS, beads=BEADS)


def test_moments_match_the_standard_normal(stats):
    n = stats["count"]
    assert n == 3 * BEADS * STEPS
    m1, m2, m3, m4, m6 = stats["sum"] / n
    # var(z^k) = E z^2k - (E z^k)^2 with E z^2 = 1, z^4 = 3, z^6 = 15, z^8 = 105, z^12 = 10395
    assert abs(m1) <= 5 * math.sqrt(1.0 / n)
    assert abs(m2 - 1.0) <= 5 * math.sqrt(2.0 / n)
    assert abs(m3) <= 5 * math.sqrt(15.0 / n)
    assert abs(m4 - 3.0) <= 5 * math.sqrt((105.0 - 9.0) / n)
    assert abs(m6 - 15.0) <= 5 * math.sqrt((10395.0 - 225.0) / n)


def test_tails_carry_the_gaussian_mass_to_six_sigma(stats):
    n = stats["count"]
    for k, got in zip(range(1, 8), stats["tail"]):
        want = n * gauss_tail(k)
        band = 5 * math.sqrt(want) + 3
        assert abs(got - want) <= band, (k, got, want)
    # 3e8 draws expect 0.6 values beyond 6 sigma and none beyond 7; the generator itself reaches 7.54
    assert 5.3 < stats["max_abs"] < 7.6


def test_components_and_steps_are_uncorrelated(stats):
    n, t = stats["count"], stats["triples"]
    assert abs(stats["lag1"]) <= 5 * math.sqrt(n)          # same bead, consecutive steps
    for c in stats["cross"]:                                 # the three normals of one bead and step
        assert abs(c) <= 5 * math.sqrt(t)
    assert abs(stats["sq_cross"] / t - 1.0) <= 5 * math.sqrt(8.0 / t)  # var(z0^2 z1^2) = 9 - 1


def test_streams_differ_by_seed_replica_and_step(engine):
    a = engine.noise_statistics(seed=1, first_step=1, steps=1, beads=4096)
    for kw in (dict(seed=2, first_step=1), dict(seed=1, first_step=2), dict(seed=1, first_step=1, replica=1)):
        b = engine.noise_statistics(steps=1, beads=4096, **kw)
        assert not np.array_equal(a["sum"], b["sum"])
    again = engine.noise_statistics(seed=1, first_step=1, steps=1, beads=4096)
    assert np.array_equal(a["tail"], again["tail"]) and a["max_abs"] == again["max_abs"]
