"""The CPU oracle against golden vectors generated from the unmodified reference
(tests/golden/make_golden.py).  Runs anywhere (no /root/reference needed).  CPU only."""
import os

import numpy as np
import pytest

import bdb200 as bd
from cases import GOLDEN, GOLDEN_SEED, GOLDEN_STEPS, build

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def test_rng_golden(oracle):
    g = np.load(os.path.join(GOLD, "rng.npz"))
    seed = int(g["seed"])
    assert np.array_equal(oracle.sfc64(seed, 4096), g["sfc64"])
    z = oracle.normals(seed, 1_000_000)
    assert np.array_equal(z[:4096], g["normals"])
    assert z.sum() == g["normals_sum"] and (z * z).sum() == g["normals_sumsq"]
    assert np.bitwise_xor.reduce(z.view(np.uint64)) == g["normals_xor"]


@pytest.mark.parametrize("tag,n", [("t100", 100), ("t1000", 1000)])
@pytest.mark.parametrize("kname,kind,period", [("open", bd.BOX_OPEN, None), ("per", bd.BOX_PERIODIC, (0.9, 1.0, 1.1))])
def test_nsearch_golden(oracle, tag, n, kname, kind, period):
    g = np.load(os.path.join(GOLD, "nsearch.npz"))
    x = g[f"{tag}_x"]
    assert np.array_equal(oracle.nsearch_buckets(kind, period, n, 0.3, x), g[f"{tag}_{kname}_buckets"])
    assert np.array_equal(oracle.nsearch_pairs(kind, period, n, 0.3, x), g[f"{tag}_{kname}_pairs"])


@pytest.mark.parametrize("name", sorted(GOLDEN))
def test_scenario_golden(oracle, name):
    g = np.load(os.path.join(GOLD, f"{name}.npz"))
    scen = GOLDEN[name]()
    assert np.array_equal(scen.positions[0], g["positions0"]), "scenario generator drifted from the fixture"
    s, ids = build(oracle, scen)
    assert np.array_equal(s.compute_forces(), g["forces"])
    assert s.compute_energy() == g["energy"]
    assert np.array_equal(s.get_neighbor_list(), g["neighbor_list"])
    for cls, key in ((bd.FF_PAIR, "pair"), (bd.FF_BOND, "bond"), (bd.FF_BEND, "bend"),
                     (bd.FF_WALL, "wall"), (bd.FF_POINT, "point")):
        for ff in ids[key]:
            assert np.array_equal(s.compute_forces(0, cls, ff), g[f"forces_{key}{ff}"])
            assert s.compute_energy(0, cls, ff) == g[f"energy_{key}{ff}"]
            if cls == bd.FF_WALL:
                assert np.array_equal(s.get_wall_stats(ff), g[f"stats_{key}{ff}"])
    assert np.array_equal(oracle.normals(GOLDEN_SEED, GOLDEN_STEPS * scen.n * 3), g["normals"])
    s.step(GOLDEN_STEPS, scen.timestep, scen.temperature, seed=GOLDEN_SEED)
    assert np.array_equal(s.get_positions(), g["positions_after"])
    assert s.compute_energy() == g["energy_after"]


# ---- the sizes BASELINE.json names, against full-size fixtures of the reference --------------------------

from cases import NAMED, NAMED_SEED, NAMED_STEPS, digest, pair_digest, sample_beads  # noqa: E402


@pytest.mark.parametrize("name", sorted(NAMED))
def test_named_size_golden(oracle, name):
    """4 000, 640, 62 600 and 10^6 beads: the oracle reproduces the reference's full arrays bit for bit
    (sha256), its Verlet pair set and a 10-step trajectory (tests/golden/make_named_golden.py)."""
    g = np.load(os.path.join(GOLD, "named_sizes.npz"))
    scen = NAMED[name]()
    assert scen.n == int(g[f"{name}.n"])
    assert digest(scen.positions[0]) == str(g[f"{name}.positions_sha"]), "scenario generator drifted from the fixture"
    s, _ = build(oracle, scen)
    f = s.compute_forces()
    assert digest(f) == str(g[f"{name}.forces_sha"])
    assert np.array_equal(f[sample_beads(scen.n)], g[f"{name}.forces_sample"])
    assert s.compute_energy() == float(g[f"{name}.energy"])
    assert np.array_equal(pair_digest(s.get_neighbor_list()), g[f"{name}.pairs"])
    s.step(NAMED_STEPS, scen.timestep, scen.temperature, seed=NAMED_SEED)
    assert digest(s.get_positions()) == str(g[f"{name}.after_sha"])
