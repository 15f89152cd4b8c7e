"""Parity of the B200 engine (libbd_b200.so, through the C ABI) with the CPU oracle and with the
golden vectors generated from the unmodified reference.  GPU only.

Tolerances (stated, not assumed):
* bucket indices, neighbour pair sets, Verlet-list pair sets, rebuild counts: bit-exact / set-equal
* forces: max_i |f_i - f_ref_i| / max_i |f_ref_i| <= 1e-12 (fp64 everywhere; only the summation
  order and FMA contraction differ); BASELINE.json asks for 1e-6
* energies, wall statistics: 1e-12 relative
* noise-replay trajectories: max |dx| <= 1e-10 over 25-30 steps (positions are O(1..10))
"""
import os

import numpy as np
import pytest

import bdb200 as bd
import scenarios as sc
from cases import GOLDEN, GOLDEN_SEED, GOLDEN_STEPS, SMALL, build, rel_force_error, sort_pairs, brute_pairs

pytestmark = pytest.mark.gpu

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
FORCE_TOL = 1e-12
ENERGY_TOL = 1e-12
TRAJ_TOL = 1e-10

CLASSES = ((bd.FF_PAIR, "pair"), (bd.FF_BOND, "bond"), (bd.FF_BEND, "bend"), (bd.FF_WALL, "wall"),
           (bd.FF_POINT, "point"))


def assert_close(a, b, tol):
    a, b = np.asarray(a, dtype=np.float64), np.asarray(b, dtype=np.float64)
    scale = max(np.abs(b).max(), 1e-300) if b.size else 1.0
    assert np.abs(a - b).max() <= tol * scale if a.size else True, (np.abs(a - b).max() / scale)


# ---- neighbour search -------------------------------------------------------------------------

@pytest.mark.parametrize("kind,period", [(bd.BOX_OPEN, None), (bd.BOX_PERIODIC, (0.9, 1.0, 1.1)),
                                         (bd.BOX_PERIODIC, (0.5, 2.0, 0.31))])
@pytest.mark.parametrize("n", [0, 1, 2, 100, 1000, 20000])
def test_buckets_bit_exact(engine, oracle, kind, period, n):
    x = np.random.default_rng(n + 7).normal(size=(n, 3)) * 1.3
    for spacing in (0.3, 0.45):
        assert np.array_equal(engine.nsearch_buckets(kind, period, n, spacing, x),
                              oracle.nsearch_buckets(kind, period, n, spacing, x))


@pytest.mark.parametrize("kind,period", [(bd.BOX_OPEN, None), (bd.BOX_PERIODIC, (0.9, 1.0, 1.1)),
                                         (bd.BOX_PERIODIC, (0.5, 2.0, 0.31))])
@pytest.mark.parametrize("n", [0, 1, 2, 100, 1000, 5000])
def test_pair_sets_equal(engine, oracle, kind, period, n):
    x = np.random.default_rng(n + 7).normal(size=(n, 3)) * 1.3
    for dcut in (0.3, 0.45):
        got = sort_pairs(engine.nsearch_pairs(kind, period, n, dcut, x))
        want = sort_pairs(oracle.nsearch_pairs(kind, period, n, dcut, x))
        assert np.array_equal(got, want)
        if n <= 1000:
            assert np.array_equal(got, sort_pairs(brute_pairs(x, dcut, period)))


@pytest.mark.parametrize("tag,n", [("t100", 100), ("t1000", 1000)])
@pytest.mark.parametrize("kname,kind,period", [("open", bd.BOX_OPEN, None), ("per", bd.BOX_PERIODIC, (0.9, 1.0, 1.1))])
def test_nsearch_golden(engine, tag, n, kname, kind, period):
    g = np.load(os.path.join(GOLD, "nsearch.npz"))
    x = g[f"{tag}_x"]
    assert np.array_equal(engine.nsearch_buckets(kind, period, n, 0.3, x), g[f"{tag}_{kname}_buckets"])
    assert np.array_equal(sort_pairs(engine.nsearch_pairs(kind, period, n, 0.3, x)),
                          sort_pairs(g[f"{tag}_{kname}_pairs"]))


def test_pair_search_dense_cluster_grows_rows(engine, oracle):
    """every point within dcut of every other: rows longer than the default capacity"""
    x = np.random.default_rng(3).normal(size=(300, 3)) * 0.01
    got = sort_pairs(engine.nsearch_pairs(bd.BOX_OPEN, None, 300, 0.3, x))
    assert len(got) == 300 * 299 // 2
    assert np.array_equal(got, sort_pairs(oracle.nsearch_pairs(bd.BOX_OPEN, None, 300, 0.3, x)))


def test_coincident_and_far_points(engine, oracle):
    x = np.zeros((50, 3))
    x[25:] += 1000.0
    got = sort_pairs(engine.nsearch_pairs(bd.BOX_OPEN, None, 50, 0.3, x))
    assert np.array_equal(got, sort_pairs(oracle.nsearch_pairs(bd.BOX_OPEN, None, 50, 0.3, x)))
    assert len(got) == 2 * (25 * 24 // 2)


# ---- forces, energies, statistics ---------------------------------------------------------------

@pytest.mark.parametrize("name", sorted(SMALL))
def test_forces_and_energies_match_oracle(engine, oracle, name):
    scen = SMALL[name]()
    se, ids = build(engine, scen)
    so, _ = build(oracle, scen)
    ftot = so.compute_forces()
    assert rel_force_error(se.compute_forces(), ftot) <= FORCE_TOL
    etot = so.compute_energy()
    assert se.compute_energy() == pytest.approx(etot, rel=ENERGY_TOL)
    total_scale = np.linalg.norm(ftot, axis=1).max()
    for cls, key in CLASSES:
        for ff in ids[key]:
            fo = so.compute_forces(0, cls, ff)
            # a forcefield whose forces cancel to rounding noise (springs at rest length) is
            # measured against the scale of the total force
            scale = max(np.linalg.norm(fo, axis=1).max(), 1e-3 * total_scale)
            assert np.linalg.norm(se.compute_forces(0, cls, ff) - fo, axis=1).max() <= FORCE_TOL * scale, (key, ff)
            assert se.compute_energy(0, cls, ff) == pytest.approx(so.compute_energy(0, cls, ff), rel=ENERGY_TOL, abs=ENERGY_TOL * abs(etot))
            if cls == bd.FF_WALL:
                assert_close(se.get_wall_stats(ff), so.get_wall_stats(ff), 1e-11)
    assert np.array_equal(sort_pairs(se.get_neighbor_list()), sort_pairs(so.get_neighbor_list()))


@pytest.mark.parametrize("name", sorted(GOLDEN))
def test_scenario_golden(engine, name):
    """against arrays produced by the unmodified reference (tests/golden/make_golden.py)"""
    g = np.load(os.path.join(GOLD, f"{name}.npz"))
    scen = GOLDEN[name]()
    assert np.array_equal(scen.positions[0], g["positions0"])
    s, ids = build(engine, scen)
    assert rel_force_error(s.compute_forces(), g["forces"]) <= FORCE_TOL
    assert s.compute_energy() == pytest.approx(float(g["energy"]), rel=ENERGY_TOL)
    assert np.array_equal(sort_pairs(s.get_neighbor_list()), sort_pairs(g["neighbor_list"]))
    total_scale = np.linalg.norm(g["forces"], axis=1).max()
    for cls, key in CLASSES:
        for ff in ids[key]:
            fo = g[f"forces_{key}{ff}"]
            scale = max(np.linalg.norm(fo, axis=1).max(), 1e-3 * total_scale)
            assert np.linalg.norm(s.compute_forces(0, cls, ff) - fo, axis=1).max() <= FORCE_TOL * scale, (key, ff)
            assert s.compute_energy(0, cls, ff) == pytest.approx(float(g[f"energy_{key}{ff}"]), rel=ENERGY_TOL, abs=ENERGY_TOL * abs(float(g["energy"])))
            if cls == bd.FF_WALL:
                assert_close(s.get_wall_stats(ff), g[f"stats_{key}{ff}"], 1e-11)
    # validation mode: replay the normals the reference drew
    s.step(GOLDEN_STEPS, scen.timestep, scen.temperature, replay_normals=g["normals"])
    assert np.abs(s.get_positions() - g["positions_after"]).max() <= TRAJ_TOL
    assert s.compute_energy() == pytest.approx(float(g["energy_after"]), rel=1e-9)


# ---- trajectories -------------------------------------------------------------------------------

@pytest.mark.parametrize("name", sorted(SMALL))
def test_replayed_trajectory_matches_oracle(engine, oracle, name):
    scen = SMALL[name]()
    se, _ = build(engine, scen)
    so, _ = build(oracle, scen)
    z = oracle.normals(5, 30 * scen.n * 3).reshape(30, scen.n, 3)
    # two chunks: the second continues the previous-noise state
    for s in (se, so):
        s.step(20, scen.timestep, scen.temperature, replay_normals=z[:20], first_step=1)
        s.step(10, scen.timestep, scen.temperature, replay_normals=z[20:], first_step=21)
    assert np.abs(se.get_positions() - so.get_positions()).max() <= TRAJ_TOL


def test_external_forces_enter_every_whole_system_evaluation(engine, oracle):
    """bd_set_external_forces: host-evaluated user forcefields (3-sim-1kb glue / loop)."""
    scen = SMALL["chromatin_1kb"]()
    se, _ = build(engine, scen)
    so, _ = build(oracle, scen)
    ext = np.random.default_rng(2).normal(size=(scen.n, 3)) * 5.0
    for s in (se, so):
        s.set_external_forces(ext)
    assert rel_force_error(se.compute_forces(), so.compute_forces()) <= 1e-12
    assert rel_force_error(se.compute_forces(ff_class=bd.FF_PAIR, ff=0), so.compute_forces(ff_class=bd.FF_PAIR, ff=0)) <= 1e-12
    z = oracle.normals(3, 12 * scen.n * 3)
    for s in (se, so):
        s.step(12, scen.timestep, scen.temperature, replay_normals=z)
    assert np.abs(se.get_positions() - so.get_positions()).max() <= TRAJ_TOL
    for s in (se, so):
        s.set_external_forces(None)
    assert rel_force_error(se.compute_forces(), so.compute_forces()) <= 1e-12


@pytest.mark.parametrize("name", ["ab_box", "ab_sphere"])
def test_long_replay_rebuilds_with_the_oracle(engine, oracle, name):
    """200 steps at a large timestep: the Verlet list is rebuilt many times.  In a periodic box the
    engine rebuilds at exactly the steps the reference does (same fp64 criterion).  In an open box
    the reference rebuilds at EVERY step -- its consistency check compares the caller's
    open_box{particle_count = 1000} with the hinted copy {particle_count = N}
    (MD/forcefield/detail/neighbor_list.hpp:37-40,114-118) -- which the engine does not imitate."""
    scen = sc.ab_box(20, 20) if name == "ab_box" else sc.ab_sphere(20, 20)
    se, _ = build(engine, scen)
    so, _ = build(oracle, scen)
    steps, dt = 200, 1e-4
    z = oracle.normals(9, steps * scen.n * 3)
    se.step(steps, dt, scen.temperature, replay_normals=z)
    so.step(steps, dt, scen.temperature, replay_normals=z)
    assert np.abs(se.get_positions() - so.get_positions()).max() <= 1e-8
    ce, co = se.get_counters(), so.get_counters()
    assert ce.steps_done == steps
    assert 3 <= ce.list_rebuilds < steps
    if name == "ab_box":
        assert ce.list_rebuilds == co.list_rebuilds
    else:
        assert co.list_rebuilds == steps


def test_adaptive_timestep_matches_oracle(engine, oracle):
    scen = SMALL["genome_10mb"]()
    se, _ = build(engine, scen)
    so, _ = build(oracle, scen)
    z = oracle.normals(9, 10 * scen.n * 3)
    for s in (se, so):
        s.step(10, 1e-3, scen.temperature, spacestep=0.01, replay_normals=z)
    assert se.get_last_timestep() == pytest.approx(so.get_last_timestep(), rel=1e-9)
    assert np.abs(se.get_positions() - so.get_positions()).max() <= TRAJ_TOL


def test_parameter_updates_every_step(engine, oracle):
    """bead_scale / bond_scale / wall semiaxes change every step in the 100kb driver
    (5-sim-genome/src/simulation_interphase/simulation_driver_interphase.cc:42-80)."""
    scen = SMALL["genome_100kb"]()
    se, ids = build(engine, scen)
    so, _ = build(oracle, scen)
    semi = np.array(scen.params["semiaxes"])
    z = oracle.normals(3, 8 * scen.n * 3).reshape(8, scen.n, 3)
    for k in range(8):
        t = sc.c4_tables(scen.params, bead_scale=0.8 + 0.02 * k, bond_scale=0.9 + 0.01 * k)
        semi = semi * 0.999
        for s in (se, so):
            s.update_pair_ff(ids["pair"][0], t["pair"], t["dcut"])
            s.update_bond_ff(ids["bond"][0], t["chain_bond"])
            s.update_bond_ff(ids["bond"][1], t["loop_bond"])
            s.update_bond_ff(ids["bond"][2], t["nuc_bond"])
            s.update_wall_ff(ids["wall"][0], (0, 0, 0), semi, t["wall_in"])
            s.update_wall_ff(ids["wall"][1], (0, 0, 0), semi)
            s.step(1, scen.timestep, scen.temperature, first_step=k + 1, replay_normals=z[k])
        assert_close(se.get_wall_stats(ids["wall"][0]), so.get_wall_stats(ids["wall"][0]), 1e-10)
    assert np.abs(se.get_positions() - so.get_positions()).max() <= TRAJ_TOL


def test_enable_disable_forcefields(engine, oracle):
    """forcefields are added/removed between phases (simulation_spindle/simulation_driver.cc:207,238)"""
    scen = SMALL["genome_10mb"]()
    se, ids = build(engine, scen)
    so, _ = build(oracle, scen)
    for s in (se, so):
        s.set_ff_enabled(bd.FF_POINT, ids["point"][1], False)
        s.set_ff_enabled(bd.FF_BEND, ids["bend"][0], False)
    assert rel_force_error(se.compute_forces(), so.compute_forces()) <= FORCE_TOL
    for s in (se, so):
        s.set_ff_enabled(bd.FF_POINT, ids["point"][1], True)
        s.set_ff_enabled(bd.FF_PAIR, ids["pair"][0], False)
    assert rel_force_error(se.compute_forces(), so.compute_forces()) <= FORCE_TOL
    assert se.compute_energy() == pytest.approx(so.compute_energy(), rel=ENERGY_TOL)


# ---- replicas -----------------------------------------------------------------------------------

def test_replica_ensemble_equals_independent_systems(engine, oracle):
    R = 5
    scen = sc.genome_10mb(total=320, replicas=R, bend_energy=0.7)
    ens = bd.System(engine, scen.n, R)
    scen.apply(ens)
    steps = 12
    z = np.stack([oracle.normals(100 + r, steps * scen.n * 3).reshape(steps, scen.n, 3) for r in range(R)], axis=1)
    ens.step(steps, scen.timestep, scen.temperature, replay_normals=z)
    for r in range(R):
        so = bd.System(oracle, scen.n)
        scen.apply(so, replicas=[r])
        f0 = so.compute_forces()
        so.step(steps, scen.timestep, scen.temperature, replay_normals=z[:, r])
        assert np.abs(ens.get_positions(r) - so.get_positions()).max() <= TRAJ_TOL, r
        assert ens.compute_energy(r) == pytest.approx(so.compute_energy(), rel=1e-9)
        del f0


# ---- production noise (Philox) -------------------------------------------------------------------

def test_zero_temperature_freezes_free_particles(engine):
    # micromd tests/simulation/test_brownian_dynamics.cc:42-57
    n = 100
    x = np.random.default_rng(0).normal(size=(n, 3))
    s = bd.System(engine, n)
    s.set_types(np.zeros(n, np.uint8), [1.0])
    s.set_positions(x)
    s.step(50, 1e-3, temperature=0.0, seed=1)
    assert np.array_equal(s.get_positions(), x)


def test_philox_is_reproducible_and_chunking_invariant(engine):
    scen = SMALL["ab_sphere"]()
    a, _ = build(engine, scen)
    b, _ = build(engine, scen)
    a.step(30, scen.timestep, scen.temperature, seed=42)
    b.step(10, scen.timestep, scen.temperature, seed=42, first_step=1)
    b.step(20, scen.timestep, scen.temperature, seed=42, first_step=11)
    assert np.array_equal(a.get_positions(), b.get_positions())
    c, _ = build(engine, scen)
    c.step(30, scen.timestep, scen.temperature, seed=43)
    assert not np.array_equal(a.get_positions(), c.get_positions())


def test_free_diffusion_statistics(engine):
    """<dx^2> = 2 kT mu t per axis for the BAOAB-limit scheme after many steps (the half-sum of
    consecutive noises loses half a step's variance at each end)."""
    n, steps, dt, kT, mu = 20000, 200, 1e-3, 0.7, 1.3
    s = bd.System(engine, n)
    s.set_types(np.zeros(n, np.uint8), [mu])
    s.set_positions(np.zeros((n, 3)))
    s.step(steps, dt, temperature=kT, seed=11)
    x = s.get_positions()
    want = 2 * kT * mu * dt * (steps - 0.5)
    got = (x * x).mean(axis=0)
    assert np.allclose(got, want, rtol=0.04), (got, want)
    assert np.abs(x.mean(axis=0)).max() < 4 * np.sqrt(want / n)


def test_harmonic_well_canonical_distribution(engine):
    """micromd tests/simulation/test_brownian_dynamics.cc:135-215: radial distribution of
    particles in a harmonic well follows the canonical distribution (KS test)."""
    from scipy import stats

    n, K, kT = 5000, 2.0, 0.8
    s = bd.System(engine, n)
    s.set_types(np.zeros(n, np.uint8), [1.0])
    s.set_positions(np.zeros((n, 3)))
    s.add_point_ff(bd.table([[bd.harmonic(K)]]), (0, 0, 0))
    s.step(4000, 2e-3, temperature=kT, seed=5)
    x = s.get_positions()
    # each coordinate ~ N(0, kT/K)
    for d in range(3):
        p = stats.kstest(x[:, d] / np.sqrt(kT / K), "norm").pvalue
        assert p > 1e-3, (d, p)


# ---- errors -------------------------------------------------------------------------------------

def test_error_reporting(engine):
    s = bd.System(engine, 10)
    with pytest.raises(bd.BackendError):
        s.set_types(np.full(10, 3, np.uint8), [1.0])
    with pytest.raises(bd.BackendError):
        s.add_bond_ff(bd.table([[bd.harmonic(1.0)]]), [[0, 10]])
    with pytest.raises(bd.BackendError):
        s.compute_forces(0, bd.FF_WALL, 0)
    with pytest.raises(bd.BackendError):
        s.step(1, -1.0)


def test_empty_system(engine):
    s = bd.System(engine, 0)
    s.step(3, 1e-3)
    assert s.get_positions().shape == (0, 3)
    assert s.compute_energy() == 0.0


def ensemble_statistics_agree(engine, oracle, scen, steps, r_gpu, r_cpu, contact, moved_observable):
    """Long runs cannot be compared bead by bead (device Philox vs the reference's sfc64 + ziggurat), so
    the ensemble statistics are: radius of gyration per chain, the number of contacts, and the CONTACT MAP
    -- contacts binned into 16 x 16 blocks of n / 16 beads, the observable the paper compares with Hi-C -- averaged
    over replicas that start from the same configuration.  Every block of the map
    must agree within 5 standard errors and the blocks together must scatter like chi-square."""
    import ensemble as en

    nbin, maps = 16, {}

    def observables(xyz, tag=None):
        rg = en.radius_of_gyration(xyz, scen.chains).mean()
        pairs = oracle.nsearch_pairs(bd.BOX_OPEN, None, scen.n, contact, xyz)
        if tag is not None and len(pairs):
            lo = np.minimum(pairs[:, 0], pairs[:, 1]) * nbin // scen.n
            hi = np.maximum(pairs[:, 0], pairs[:, 1]) * nbin // scen.n
            maps.setdefault(tag, []).append(np.bincount(lo * nbin + hi, minlength=nbin * nbin).astype(np.float64))
        return rg, len(pairs)

    gpu = en.ReplicaEnsemble(engine, scen, r_gpu, seed=21)
    gpu.step(steps)
    a = np.array([observables(x, "gpu") for x in gpu.positions()])
    cpu = en.ReplicaEnsemble(oracle, scen, r_cpu, seed=5)
    cpu.step(steps)
    b = np.array([observables(x, "cpu") for x in cpu.positions()])
    start = np.array(observables(scen.positions[0]))
    for k, name in enumerate(("radius of gyration", "contacts")):
        err = np.sqrt(a[:, k].var(ddof=1) / r_gpu + b[:, k].var(ddof=1) / r_cpu)
        assert abs(a[:, k].mean() - b[:, k].mean()) < 4.5 * err, (name, a[:, k].mean(), b[:, k].mean(), err)
    # the binned contact map, block by block
    ma, mb = np.array(maps["gpu"]), np.array(maps["cpu"])
    se = np.sqrt(ma.var(axis=0, ddof=1) / r_gpu + mb.var(axis=0, ddof=1) / r_cpu)
    used = se > 0
    assert used.sum() >= nbin  # at least the diagonal blocks carry contacts
    zscore = (ma.mean(axis=0) - mb.mean(axis=0))[used] / se[used]
    assert np.abs(zscore).max() < 5.0, (np.abs(zscore).max(), int(used.sum()))
    assert abs((zscore ** 2).sum() - used.sum()) < 6.0 * np.sqrt(2.0 * used.sum()) + 0.15 * used.sum(), (zscore ** 2).sum()
    # the run did move the chains away from their start by far more than the resolution of the test
    k = moved_observable
    assert abs(b[:, k].mean() - start[k]) > 20 * np.sqrt(b[:, k].var(ddof=1) / r_cpu)
    gpu.close()
    cpu.close()


def test_statistical_equivalence_with_reference_noise(engine, oracle):
    """the 10Mb ana/telophase model (320 beads, blocks of 20 beads), 1500 steps"""
    ensemble_statistics_agree(engine, oracle, sc.genome_10mb(total=320, bend_energy=0.7), steps=1500, r_gpu=192, r_cpu=48,
                              contact=0.25, moved_observable=0)


@pytest.mark.slow
def test_statistical_equivalence_on_the_100kb_model(engine, oracle):
    """the 100kb interphase model -- the one the paper compares with Hi-C -- at 3 060 beads in its ellipsoid (46 chains,
    nucleoli, both walls): 400 steps take the overlapping start to a quarter fewer contacts; radius of gyration,
    contact count and the 16 x 16 contact map of 96 device replicas against 16 replicas of the CPU restatement with
    the reference's own generator"""
    ensemble_statistics_agree(engine, oracle, sc.genome_100kb(3000, 30), steps=400, r_gpu=96, r_cpu=16, contact=0.36,
                              moved_observable=1)


@pytest.mark.gpu
def test_contact_map_accumulates_on_the_device(engine, oracle):
    """bd_contact_* against contact_map.cc:25-91 restated with the oracle's neighbour searcher:
    one count per unordered pair within the contact distance per update, open box even when the
    simulation is periodic, triplets in (i, j) order, clear() empties the map."""
    from collections import Counter

    s = sc.chromatin_1kb(n_chains=6, chain_len=200)   # periodic unit cell: the map must ignore it
    e, _ = build(engine, s)
    distance = 0.24
    want = Counter()
    assert e.contact_accumulate().shape == (0, 3)
    for k in range(6):
        e.step(15, s.timestep, s.temperature, seed=5, first_step=1 + 15 * k)
        e.contact_update(distance)
        x = e.get_positions()
        pairs = oracle.nsearch_pairs(bd.BOX_OPEN, None, s.n, distance, x)
        assert np.array_equal(sort_pairs(pairs), sort_pairs(brute_pairs(x, distance)))
        want.update(map(tuple, sort_pairs(pairs).tolist()))
        if k == 2:                                   # the engine keeps stepping on the same list afterwards
            got = e.contact_accumulate()
            assert len(got) == len(want)
    got = e.contact_accumulate()
    ref = np.array([(i, j, c) for (i, j), c in sorted(want.items())], dtype=np.uint32)
    assert got.dtype == np.uint32 and np.array_equal(got, ref)
    assert (got[:, 0] < got[:, 1]).all() and got[:, 2].max() > 1
    e.contact_clear()
    assert e.contact_accumulate().shape == (0, 3)
    e.contact_update(distance * 1.5)                 # a new distance re-tunes the searcher
    x = e.get_positions()
    one = sort_pairs(brute_pairs(x, distance * 1.5))
    got = e.contact_accumulate()
    assert np.array_equal(got[:, :2], one.astype(np.uint32)) and (got[:, 2] == 1).all()
    e.close()


@pytest.mark.gpu
def test_contact_map_of_one_replica_of_an_ensemble(engine):
    s = sc.genome_10mb(total=320, replicas=3)
    e = bd.System(engine, s.n, 3)
    s.apply(e)
    e.step(20, s.timestep, s.temperature, seed=2)
    for r in (2, 0):
        e.contact_clear()
        e.contact_update(0.3, replica=r)
        got = e.contact_accumulate()
        assert np.array_equal(got[:, :2], sort_pairs(brute_pairs(e.get_positions(replica=r), 0.3)).astype(np.uint32))
    e.close()


@pytest.mark.gpu
@pytest.mark.parametrize("bits", [16, 8, -1])
def test_snapshot_floats_are_the_stores_bit_for_bit(engine, oracle, bits):
    """bd_snapshot_* against simulation_store.cc:253-276,403-407 (quantize) restated in the oracle
    and in numpy; the engine keeps stepping while the copy is in flight."""
    s = sc.genome_10mb(total=320, replicas=3)
    e = bd.System(engine, s.n, 3)
    s.apply(e)
    x = np.stack([e.get_positions(r) for r in range(3)])
    # ties of the rounding, negative values, values beyond float's integer range
    wild = x[0].copy()
    wild[:6] = [[0.5 / 65536, 1.5 / 65536, 2.5 / 65536], [-0.5 / 65536, -1.5 / 65536, -2.5 / 65536],
                [1e9 + 0.3, -1e9 - 0.3, 3.0000076], [123.4567891234, -0.0, 0.0],
                [1 + 2.0 ** -24, 1 + 2.0 ** -23, 1 - 2.0 ** -25], [2.0 ** -30, -2.0 ** -30, 1e-45]]

    def want(a):
        a = np.asarray(a).reshape(-1, 3)
        ref = oracle.quantize_positions(a, bits)
        f = a.astype(np.float32)
        mine = f if bits < 0 else np.rint(f * np.float32(1 << bits)) / np.float32(1 << bits)
        assert mine.dtype == np.float32 and np.array_equal(ref.view(np.uint32), mine.view(np.uint32))
        return ref

    e.set_positions(wild, replica=1)
    e.snapshot_begin(replica=1, fraction_bits=bits)
    got = e.snapshot_wait()
    assert got.dtype == np.float32 and np.array_equal(got.view(np.uint32), want(wild).view(np.uint32))
    e.set_positions(x[1], replica=1)
    e.snapshot_begin(replica=-1, fraction_bits=bits)
    e.step(5, s.timestep, s.temperature, seed=9)     # overlaps the copy; the snapshot is of the state before
    got = e.snapshot_wait()
    assert np.array_equal(got.view(np.uint32), want(x).view(np.uint32))
    after = np.stack([e.get_positions(r) for r in range(3)])
    assert not np.array_equal(after, x)
    e.snapshot_begin(replica=-1, fraction_bits=bits)
    assert np.array_equal(e.snapshot_wait().view(np.uint32), want(after).view(np.uint32))
    with pytest.raises(bd.BackendError):
        e.snapshot_wait()
    e.close()


@pytest.mark.gpu
def test_dynamic_pair_lists_on_the_device(engine, oracle):
    """glue- and loop-shaped pair lists edited between steps (set_bond_pairs), the glue
    list through the periodic box (set_bond_min_image); engine against the oracle, which
    tests/test_oracle_vs_reference.py pins to micromd for the same calls."""
    rng = np.random.default_rng(17)
    s = sc.chromatin_1kb(n_chains=3, chain_len=120)
    L, n = np.asarray(s.period), s.n
    well = bd.pair_table(s.ntypes, lambda ti, tj: [bd.softwell(6, -1.5, 0.4)])
    spring = bd.pair_table(s.ntypes, lambda ti, tj: [bd.spring(40.0, 0.3)])

    def random_pairs(m):
        i = rng.integers(0, n, size=m)
        j = (i + rng.integers(1, n - 1, size=m)) % n
        return np.stack([i, j], axis=1).astype(np.uint32)

    x = s.positions[0] + np.random.default_rng(5).integers(-2, 3, size=(n, 3)) * L   # periodic images
    systems, ids = [], []
    for backend in (engine, oracle):
        e, _ = build(backend, s)
        e.set_positions(x)
        glue = e.add_bond_ff(well, np.zeros((0, 2), np.uint32))
        loop = e.add_bond_ff(spring, np.zeros((0, 2), np.uint32))
        e.set_bond_min_image(glue, True)
        systems.append(e)
        ids.append((glue, loop))
    z = oracle.normals(3, 6 * 4 * n * 3).reshape(6, 4, n, 3)
    for round_ in range(6):
        gp, lp = random_pairs(40 + 5 * round_), random_pairs(12)
        if round_ == 4:
            gp = gp[:0]
        out = []
        for e, (glue, loop) in zip(systems, ids):
            e.set_bond_pairs(glue, gp)
            e.set_bond_pairs(loop, lp)
            out.append((e.compute_forces(ff_class=bd.FF_BOND, ff=glue), e.compute_forces(ff_class=bd.FF_BOND, ff=loop),
                        e.compute_energy(ff_class=bd.FF_BOND, ff=glue), e.compute_forces(), e.compute_energy()))
            e.step(4, s.timestep, s.temperature, replay_normals=z[round_])
        (fg, fl, eg, fa, ea), (og, ol, oeg, oa, oea) = out
        assert rel_force_error(fl, ol) <= FORCE_TOL and rel_force_error(fa, oa) <= FORCE_TOL
        if len(gp):
            assert np.abs(og).max() > 0 and rel_force_error(fg, og) <= FORCE_TOL
        else:
            assert not fg.any()
        assert eg == pytest.approx(oeg, rel=1e-9, abs=1e-12) and ea == pytest.approx(oea, rel=1e-9)
        assert np.abs(systems[0].get_positions() - systems[1].get_positions()).max() <= TRAJ_TOL
    with pytest.raises(bd.BackendError):
        systems[0].set_bond_pairs(ids[0][0], np.array([[0, n]], np.uint32))
    for e in systems:
        e.close()


@pytest.mark.gpu
def test_replacing_a_pair_list_does_not_rebuild_the_verlet_list(engine):
    """bd_set_bond_pairs refreshes the slot-space bond rows only (k_bond_rows): the number of list builds
    stays what the displacement criterion asks for, and the forces are those of a fresh engine."""
    s = sc.chromatin_1kb(n_chains=4, chain_len=150)
    n = s.n
    spring = bd.pair_table(s.ntypes, lambda ti, tj: [bd.spring(40.0, 0.3)])
    rng = np.random.default_rng(2)
    e, _ = build(engine, s)
    ff = e.add_bond_ff(spring, np.zeros((0, 2), np.uint32))
    e.step(3, s.timestep, s.temperature, seed=1)
    before = e.get_counters().list_rebuilds
    for k in range(10):
        i = rng.integers(0, n - 40, size=20)
        pairs = np.stack([i, i + rng.integers(2, 40, size=20)], axis=1).astype(np.uint32)
        e.set_bond_pairs(ff, pairs)
        f = e.compute_forces()
    assert e.get_counters().list_rebuilds == before
    fresh, _ = build(engine, s)
    ff2 = fresh.add_bond_ff(spring, pairs)
    fresh.set_positions(e.get_positions())
    assert rel_force_error(f, fresh.compute_forces()) <= FORCE_TOL
    assert ff2 == ff
    e.close()
    fresh.close()


@pytest.mark.gpu
def test_pair_search_and_distances_on_resident_positions(engine, oracle):
    """bd_search_pairs / bd_pair_distances: what glue_simulator.cpp:35-80 asks of the positions, answered
    on the device -- pair sets equal to the reference searcher's on the same coordinates, in sorted order."""
    s = sc.chromatin_1kb(n_chains=5, chain_len=160)
    L = np.asarray(s.period)
    e, _ = build(engine, s)
    for k in range(3):
        e.step(12, s.timestep, s.temperature, seed=4, first_step=1 + 12 * k)
        x = e.get_positions()
        for dist in (0.33, 0.5):
            boxed = e.search_pairs(dist, use_box=True)
            want = sort_pairs(oracle.nsearch_pairs(bd.BOX_PERIODIC, L, s.n, dist, x))
            assert np.array_equal(boxed, want) and np.array_equal(want, sort_pairs(brute_pairs(x, dist, L)))
            opened = e.search_pairs(dist, use_box=False)
            assert np.array_equal(opened, sort_pairs(brute_pairs(x, dist)))
            assert len(boxed) >= len(opened) > 0
        pairs = boxed[:: max(1, len(boxed) // 50)]
        d = x[pairs[:, 0]] - x[pairs[:, 1]]
        assert np.array_equal(e.pair_distances(pairs), np.sqrt(d[:, 0] * d[:, 0] + d[:, 1] * d[:, 1] + d[:, 2] * d[:, 2]))
        dm = d - np.rint(d * (1 / L)) * L
        got = e.pair_distances(pairs, min_image=True)
        assert np.abs(got - np.sqrt((dm * dm).sum(axis=1))).max() <= 1e-15 and (got <= 0.5 * (1 + 1e-12)).all()
    # the contact map shares the searcher: switching between the two uses keeps both right
    e.contact_update(0.33)
    assert np.array_equal(e.contact_accumulate()[:, :2], sort_pairs(brute_pairs(e.get_positions(), 0.33)))
    assert e.pair_distances(np.zeros((0, 2), np.uint32)).shape == (0,)
    with pytest.raises(bd.BackendError):
        e.pair_distances(np.array([[0, s.n]], np.uint32))
    e.close()


# ---- tile staging (engine option "staged") ----------------------------------------------------------

@pytest.mark.parametrize("name", ["genome_100kb", "chromatin_1kb", "ab_sphere"])
def test_staged_pair_walk_gives_the_same_forces_and_trajectory(engine, oracle, name):
    """bd_set_option("staged", 1): Verlet rows of 16-bit positions inside the tile's staged neighbourhood,
    shadows copied to shared memory by bulk asynchronous copies (bd_device.cuh: tile staging).  Same listed
    set, same exact evaluation: forces, energy, exported list and a replayed trajectory equal the oracle's,
    for an open-box, a periodic and a wall-confined model.  (Measured slower than the global gathers on
    B200 -- profiles/README.md -- so it is off by default; this keeps the path honest.)"""
    scen = SMALL[name]()
    se, _ = build(engine, scen)
    se.set_option("staged", 1)
    so, _ = build(oracle, scen)
    assert se.get_option("staged") == 1
    assert rel_force_error(se.compute_forces(), so.compute_forces()) <= FORCE_TOL
    assert se.compute_energy() == pytest.approx(so.compute_energy(), rel=ENERGY_TOL)
    assert np.array_equal(sort_pairs(se.get_neighbor_list()), sort_pairs(so.get_neighbor_list()))
    steps = 30
    z = oracle.normals(17, steps * scen.n * 3)
    se.step(steps, scen.timestep, scen.temperature, replay_normals=z)
    so.step(steps, scen.timestep, scen.temperature, replay_normals=z)
    assert se.get_option("staged") == 1 and se.get_option("max_stage") > 0
    assert np.abs(se.get_positions() - so.get_positions()).max() <= TRAJ_TOL
    # a capacity below the largest tile: those tiles translate through the run table instead
    se.set_option("stage_capacity", 64)
    assert rel_force_error(se.compute_forces(), so.compute_forces()) <= FORCE_TOL


@pytest.mark.parametrize("kind,period", [(bd.BOX_OPEN, None), (bd.BOX_PERIODIC, (0.9, 1.0, 1.1)),
                                         (bd.BOX_PERIODIC, (0.5, 2.0, 0.31))])
@pytest.mark.parametrize("n", [2, 100, 1000, 5000])
def test_ordered_pair_search_is_the_reference_emission_order(engine, oracle, kind, period, n):
    """bd_nsearch_pairs_ordered: not only the reference's pair SET but its ORDER (bucket by bucket, directed
    neighbours ascending, neighbour-bucket member outer) -- the oracle's search emits in the reference's order
    bit for bit (tests/test_oracle_vs_reference.py::test_buckets_and_pairs_bit_exact), so array equality."""
    x = sc.random_points(n, seed=n + 3)
    got = engine.nsearch_pairs(kind, period, n, 0.3, x, ordered=True)
    want = oracle.nsearch_pairs(kind, period, n, 0.3, x)
    assert np.array_equal(got, want)


@pytest.mark.parametrize("kind,period", [(bd.BOX_OPEN, None), (bd.BOX_PERIODIC, (0.9, 1.0, 1.1))])
@pytest.mark.parametrize("n", [1, 33, 1000])
def test_batched_query_is_the_brute_force_set(engine, kind, period, n):
    """bd_nsearch_query = md::neighbor_searcher::query for every query point: strictly closer than dcut, with the
    box's shortest displacement, ascending indices -- the brute-force expectation of the reference's own unit test
    (micromd tests/misc/test_neighbor_searcher.cc), in the reference's fp64 arithmetic."""
    x = sc.random_points(n, seed=n + 1)
    queries = np.concatenate([x[: min(n, 40)], sc.random_points(25, seed=99)])
    start, index = engine.nsearch_query(kind, period, 0.3, x, queries)
    assert start[0] == 0 and start[-1] == len(index) and (np.diff(start) >= 0).all()
    for q, point in enumerate(queries):
        d = x - point
        if period is not None:
            L = np.asarray(period)
            d = d - np.rint(d * (1 / L)) * L
        r2 = d[:, 0] * d[:, 0] + d[:, 1] * d[:, 1] + d[:, 2] * d[:, 2]
        want = np.nonzero(r2 < 0.3 * 0.3)[0]
        assert np.array_equal(index[start[q]:start[q + 1]], want), q


def test_row_length_ranking_is_an_option_with_the_same_results(engine, oracle):
    """bd_set_option("balance_rows", 1): the round-1 thread order (beads of a tile ranked by row length, rows
    moved to their thread positions).  Default is slot order; both give the oracle's forces and trajectory."""
    scen = SMALL["genome_100kb"]()
    so, _ = build(oracle, scen)
    fo = so.compute_forces()
    z = oracle.normals(21, 20 * scen.n * 3)
    so.step(20, scen.timestep, scen.temperature, replay_normals=z)
    for flag in (0, 1):
        se, _ = build(engine, scen)
        se.set_option("balance_rows", flag)
        assert se.get_option("balance_rows") == flag
        assert rel_force_error(se.compute_forces(), fo) <= FORCE_TOL
        se.step(20, scen.timestep, scen.temperature, replay_normals=z)
        assert np.abs(se.get_positions() - so.get_positions()).max() <= TRAJ_TOL
        se.close()


# ---- mixed-precision pair forces (engine option "pair_precision" = 32) -----------------------------------

F32_FORCE_TOL = 5e-5   # of the largest force; measured 2e-6 .. 1e-5 on these systems (float coordinate offsets)
F32_TRAJ_TOL = 5e-6    # 30 replayed steps, positions O(1..10): 30 x mobility x timestep x the force error


@pytest.mark.parametrize("name", ["genome_100kb", "chromatin_1kb", "ab_box", "ab_sphere"])
def test_single_precision_pair_walk_stays_within_its_stated_tolerance(engine, oracle, name):
    """bd_set_option("pair_precision", 32): the production lane north_star describes ("SoA float4 positions") --
    the displacement to a listed neighbour is the difference of the 16-byte single-precision shadows, nothing
    else of the neighbour is read; potential and accumulation stay in double.  Not the reference's arithmetic:
    the stated tolerance is the rounding of a float coordinate offset (2^-24 of the box extent) through the
    stiff softcore<8,3>.  Bonded, wall and point forces stay exact; energies are always exact."""
    scen = SMALL[name]()
    se, _ = build(engine, scen)
    so, _ = build(oracle, scen)
    fo = so.compute_forces()
    exact = rel_force_error(se.compute_forces(), fo)
    se.set_option("pair_precision", 32)
    assert se.get_option("pair_precision") == 32
    err = rel_force_error(se.compute_forces(), fo)
    # only the upper bounds are claims; that the option took effect is read back above (a lower bound on the error
    # would fail on a system where single-precision offsets happen to be exact)
    assert exact <= FORCE_TOL and err <= F32_FORCE_TOL, (exact, err)
    assert se.compute_energy() == pytest.approx(so.compute_energy(), rel=ENERGY_TOL)
    z = oracle.normals(31, 30 * scen.n * 3)
    se.step(30, scen.timestep, scen.temperature, replay_normals=z)
    so.step(30, scen.timestep, scen.temperature, replay_normals=z)
    drift = np.abs(se.get_positions() - so.get_positions()).max()
    print(f"pair_precision=32 {name}: force error {err:.2e} of the largest force, 30-step drift {drift:.2e}")
    assert drift <= F32_TRAJ_TOL, drift
    se.set_option("pair_precision", 64)
    se.set_positions(so.get_positions())
    assert rel_force_error(se.compute_forces(), so.compute_forces()) <= FORCE_TOL  # back on the exact lane
