"""Per-replica wall geometry and the interphase driver's wall feedback in an ensemble.  GPU only.

The 100kb interphase driver moves the wall semiaxes of its system after every step from that system's
own wall reaction (5-sim-genome/src/simulation_interphase/simulation_driver_interphase.cc:70-80,
simulation_driver_forcefield.cc:230-234):

    net      = inward.stats.axial_reaction + outward.stats.axial_reaction - spring o semiaxes
    semiaxes += timestep * wall_mobility * net

A batched ensemble therefore needs one geometry per replica.  Checked here against R independent
oracle systems whose semiaxes diverge: (1) the host-driven loop through bd_get_wall_stats_all /
bd_update_wall_ff_replica, (2) the device-resident law (bd_set_wall_relaxation) in ONE multi-step call.
Tolerances: positions 1e-10 (TRAJ_TOL), semiaxes 1e-9 relative (a feedback loop over the reaction
statistics, whose summation order differs: slot order on the device, bead order in the reference).
"""
import numpy as np
import pytest

import bdb200 as bd
import scenarios as sc

pytestmark = pytest.mark.gpu

TRAJ_TOL = 1e-10
SEMI_TOL = 1e-9
R, STEPS = 4, 12
SPRING = np.array([1.0e4, 1.2e4, 0.9e4])
MOBILITY = 1.0e-4


def driver_update(semi, inward, outward, dt):
    """update_wall_semiaxes of the driver, one system; stats = {reaction, axial x, y, z}"""
    net = np.zeros(3)
    net += inward[1:] + outward[1:]
    net -= SPRING * semi
    return semi + dt * MOBILITY * net


def setup(engine, oracle):
    scen = sc.genome_100kb(3000, 30, seed=9, replicas=R, bead_scale=0.9, bond_scale=0.95)
    ens = bd.System(engine, scen.n, R)
    ids = scen.apply(ens)
    singles = []
    for r in range(R):
        so = bd.System(oracle, scen.n)
        scen.apply(so, replicas=[r])
        singles.append(so)
    semi0 = np.array(scen.params["semiaxes"])
    # every replica starts with its own ellipsoid (squeezed differently, so that beads press on the walls)
    semis = np.stack([semi0 * (0.86 + 0.02 * r) * np.array([1.0, 1.0 - 0.01 * r, 1.0 + 0.01 * r]) for r in range(R)])
    for r in range(R):
        for w in ids["wall"]:
            ens.update_wall_ff(w, (0, 0, 0), semis[r], replica=r)
            singles[r].update_wall_ff(w, (0, 0, 0), semis[r])
    z = np.stack([oracle.normals(500 + r, STEPS * scen.n * 3).reshape(STEPS, scen.n, 3) for r in range(R)], axis=1)
    return scen, ens, ids, singles, semis, z


def oracle_loop(scen, singles, ids, semis, z):
    semis = semis.copy()
    stats = np.zeros((R, 2, 4))
    for k in range(STEPS):
        for r, so in enumerate(singles):
            so.step(1, scen.timestep, scen.temperature, first_step=k + 1, replay_normals=z[k, r])
            stats[r, 0] = so.get_wall_stats(ids["wall"][0])
            stats[r, 1] = so.get_wall_stats(ids["wall"][1])
            semis[r] = driver_update(semis[r], stats[r, 0], stats[r, 1], scen.timestep)
            for w in ids["wall"]:
                so.update_wall_ff(w, (0, 0, 0), semis[r])
    return semis, stats


def test_per_replica_geometry_enters_forces_and_energy(engine, oracle):
    scen, ens, ids, singles, semis, _ = setup(engine, oracle)
    for r in range(R):
        c, s = ens.get_wall_geometry(ids["wall"][0], r)
        assert np.array_equal(s, semis[r]) and np.array_equal(c, np.zeros(3))
        fe, fo = ens.compute_forces(r), singles[r].compute_forces()
        assert np.linalg.norm(fe - fo, axis=1).max() <= 1e-12 * np.linalg.norm(fo, axis=1).max()
        for w in ids["wall"]:
            a, b = ens.get_wall_stats(w, r), singles[r].get_wall_stats(w)
            assert np.abs(a - b).max() <= 1e-11 * max(np.abs(b).max(), 1.0)
        assert ens.compute_energy(r) == pytest.approx(singles[r].compute_energy(), rel=1e-12)
    # the walls are touched, and differently in every replica: the test below is not vacuous
    inward = ens.get_wall_stats_all(ids["wall"][0])
    assert (inward[:, 0] > 0).all() and len(set(np.round(inward[:, 0], 6))) == R


def test_wall_feedback_host_driven_per_replica(engine, oracle):
    scen, ens, ids, singles, semis, z = setup(engine, oracle)
    want_semis, want_stats = oracle_loop(scen, singles, ids, semis, z)
    got = semis.copy()
    for k in range(STEPS):
        ens.step(1, scen.timestep, scen.temperature, first_step=k + 1, replay_normals=z[k])
        inward = ens.get_wall_stats_all(ids["wall"][0])
        outward = ens.get_wall_stats_all(ids["wall"][1])
        for r in range(R):
            got[r] = driver_update(got[r], inward[r], outward[r], scen.timestep)
            for w in ids["wall"]:
                ens.update_wall_ff(w, (0, 0, 0), got[r], replica=r)
    assert np.abs(got / want_semis - 1).max() <= SEMI_TOL
    assert np.abs(inward - want_stats[:, 0]).max() <= 1e-9 * np.abs(want_stats[:, 0]).max()
    assert np.abs(got - semis).max() > 1e-6  # the walls did move
    for r in range(R):
        assert np.abs(ens.get_positions(r) - singles[r].get_positions()).max() <= TRAJ_TOL, r


def test_wall_feedback_on_the_device_in_one_call(engine, oracle):
    scen, ens, ids, singles, semis, z = setup(engine, oracle)
    want_semis, want_stats = oracle_loop(scen, singles, ids, semis, z)
    ens.set_wall_relaxation(ids["wall"], SPRING, MOBILITY)
    launches = ens.get_launch_count()
    ens.step(STEPS, scen.timestep, scen.temperature, replay_normals=z)  # no host round trip between steps
    assert ens.get_launch_count() - launches >= 2 * STEPS
    for r in range(R):
        for w in ids["wall"]:
            _, s = ens.get_wall_geometry(w, r)
            assert np.abs(s / want_semis[r] - 1).max() <= SEMI_TOL, (r, w)
        assert np.abs(ens.get_positions(r) - singles[r].get_positions()).max() <= TRAJ_TOL, r
    # statistics of the last step are still readable, per replica
    inward = ens.get_wall_stats_all(ids["wall"][0])
    assert np.abs(inward - want_stats[:, 0]).max() <= 1e-9 * np.abs(want_stats[:, 0]).max()
    # switching the law off freezes the walls again
    ens.set_wall_relaxation([], SPRING, MOBILITY)
    before = ens.get_wall_geometry(ids["wall"][0], 1)[1]
    ens.step(3, scen.timestep, scen.temperature, seed=3, first_step=STEPS + 1)
    assert np.array_equal(ens.get_wall_geometry(ids["wall"][0], 1)[1], before)


def test_table_only_update_keeps_the_moving_geometry(engine):
    scen = sc.genome_100kb(1200, 12, seed=2, replicas=2)
    ens = bd.System(engine, scen.n, 2)
    ids = scen.apply(ens)
    ens.set_wall_relaxation(ids["wall"], SPRING, MOBILITY)
    ens.step(5, scen.timestep, scen.temperature, seed=1)
    moved = ens.get_wall_geometry(ids["wall"][0], 1)[1]
    t = sc.c4_tables(scen.params, bead_scale=0.97)
    ens.update_wall_ff(ids["wall"][0], None, None, t["wall_in"])
    assert np.array_equal(ens.get_wall_geometry(ids["wall"][0], 1)[1], moved)
    with pytest.raises(bd.BackendError):
        ens.update_wall_ff(ids["wall"][0], (0, 0, 0), (1.0, -1.0, 1.0))
