"""Engine parity at the sizes BASELINE.json names: 4 000 beads (C1, C2), 1 024 x 640 (C3), 64 x 62 600 (C4),
10^6 (C5).  GPU only.  These are the sizes at which Morton bit shedding (bd_nlist.cuh: k_make_grid), the
multi-tile scan, row-capacity regrowth and large blockIdx.y first occur; the miniatures of
tests/test_gpu_parity.py do not reach them.

Two references: the fixtures made from the UNMODIFIED micromd at full size
(tests/golden/named_sizes.npz: 1 000 sampled beads, pair-set digests, total energy) and the CPU oracle,
which reproduces those fixtures bit for bit (tests/test_oracle_golden.py) and runs here on the same inputs.

Tolerances: forces 1e-12 of the largest force (BASELINE asks 1e-6); energy 1e-11; Verlet pair sets
bit-exact; 10-step replayed trajectories 1e-10.
"""
import os

import numpy as np
import pytest

import bdb200 as bd
import scenarios as sc
from cases import (NAMED, NAMED_SEED, NAMED_STEPS, build, pair_digest, rel_force_error, sample_beads, sort_pairs)

pytestmark = pytest.mark.gpu

GOLD = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "named_sizes.npz"))
FORCE_TOL, ENERGY_TOL, TRAJ_TOL = 1e-12, 1e-11, 1e-10


@pytest.mark.parametrize("name", sorted(NAMED))
def test_named_size_against_the_reference_fixture(engine, oracle, name):
    scen = NAMED[name]()
    s, _ = build(engine, scen)
    idx = sample_beads(scen.n)
    f = s.compute_forces()
    scale = float(GOLD[f"{name}.force_scale"])
    assert np.linalg.norm(f[idx] - GOLD[f"{name}.forces_sample"], axis=1).max() <= FORCE_TOL * scale
    assert s.compute_energy() == pytest.approx(float(GOLD[f"{name}.energy"]), rel=ENERGY_TOL)
    assert np.array_equal(pair_digest(s.get_neighbor_list()), GOLD[f"{name}.pairs"])
    z = oracle.normals(NAMED_SEED, NAMED_STEPS * scen.n * 3)
    s.step(NAMED_STEPS, scen.timestep, scen.temperature, replay_normals=z)
    assert np.abs(s.get_positions()[idx] - GOLD[f"{name}.after_sample"]).max() <= TRAJ_TOL
    s.close()


def check_replica(ens, oracle, scen, src, r, ids=None):
    """forces, energy and Verlet pair set of replica r of an ensemble vs an oracle system on the same input"""
    so = bd.System(oracle, scen.n)
    scen.apply(so, replicas=[src])
    fo = so.compute_forces()
    assert rel_force_error(ens.compute_forces(r), fo) <= FORCE_TOL, r
    assert ens.compute_energy(r) == pytest.approx(so.compute_energy(), rel=ENERGY_TOL), r
    assert np.array_equal(sort_pairs(ens.get_neighbor_list(r)), sort_pairs(so.get_neighbor_list())), r
    return so


def test_c4_ensemble_of_64_replicas(engine, oracle):
    """the launch bench.py times: 64 x 62 600 beads in one engine; replicas 0, 37 and 63 against the oracle"""
    R, distinct = 64, 4
    scen = sc.genome_100kb(62000, 300, seed=4, replicas=distinct)
    ens = bd.System(engine, scen.n, R)
    scen.apply(ens, replicas=[])
    for r in range(R):
        ens.set_positions(scen.positions[r % distinct], replica=r)
    for r in (0, 37, 63):
        check_replica(ens, oracle, scen, r % distinct, r)
    c = ens.get_counters()
    assert c.overflow_beads == 0
    ens.close()


def test_c4_ten_step_replay_in_an_ensemble(engine, oracle):
    R = 6
    scen = sc.genome_100kb(62000, 300, seed=4, replicas=2)
    ens = bd.System(engine, scen.n, R)
    scen.apply(ens, replicas=[])
    for r in range(R):
        ens.set_positions(scen.positions[r % 2], replica=r)
    steps = NAMED_STEPS
    z = np.stack([oracle.normals(70 + r, steps * scen.n * 3).reshape(steps, scen.n, 3) for r in range(R)], axis=1)
    ens.step(steps, scen.timestep, scen.temperature, replay_normals=z)
    for r in (0, 5):
        so = bd.System(oracle, scen.n)
        scen.apply(so, replicas=[r % 2])
        so.step(steps, scen.timestep, scen.temperature, replay_normals=z[:, r])
        assert np.abs(ens.get_positions(r) - so.get_positions()).max() <= TRAJ_TOL, r
        assert rel_force_error(ens.compute_forces(r), so.compute_forces()) <= FORCE_TOL, r
    ens.close()


def test_c3_ensemble_of_1024_replicas(engine, oracle):
    """1 024 x 640 beads (blockIdx.y up to 1 023): replicas 0, 511 and 1 023, forces and a 10-step replay"""
    R, distinct = 1024, 8
    scen = sc.genome_10mb(640, replicas=distinct, bend_energy=0.7)
    ens = bd.System(engine, scen.n, R)
    scen.apply(ens, replicas=[])
    for r in range(R):
        ens.set_positions(scen.positions[r % distinct], replica=r)
    picks = (0, 511, 1023)
    singles = {r: check_replica(ens, oracle, scen, r % distinct, r) for r in picks}
    steps = NAMED_STEPS
    rng = np.random.default_rng(5)
    z = rng.standard_normal((steps, R, scen.n, 3))
    ens.step(steps, scen.timestep, scen.temperature, replay_normals=z)
    for r in picks:
        singles[r].step(steps, scen.timestep, scen.temperature, replay_normals=z[:, r])
        assert np.abs(ens.get_positions(r) - singles[r].get_positions()).max() <= TRAJ_TOL, r
    ens.close()


def test_million_bead_pair_set_equals_the_searcher(engine, oracle):
    """C5 at 10^6 beads: the engine's Verlet list (radius 1.5 dcut) is the pair set of the reference's
    searcher on the same points, and the standalone device searcher agrees"""
    scen = NAMED["chromatin_1kb_1M"]()
    s, _ = build(engine, scen)
    dcut = max(ff["dcut"] for ff in scen.pair_ffs)
    want = sort_pairs(oracle.nsearch_pairs(bd.BOX_PERIODIC, scen.period, scen.n, 1.5 * dcut, scen.positions[0],
                                           cap=20_000_000))
    assert np.array_equal(sort_pairs(s.get_neighbor_list()), want)
    got = sort_pairs(engine.nsearch_pairs(bd.BOX_PERIODIC, scen.period, scen.n, 1.5 * dcut, scen.positions[0],
                                          cap=20_000_000))
    assert np.array_equal(got, want)
    s.close()
