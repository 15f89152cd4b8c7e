"""Pins the CPU oracle (oracle/bd_oracle.c) to the unmodified reference (oracle/_ref, built from
/root/reference by oracle/Makefile).  Skipped where the reference library is absent; the golden
fixtures (test_oracle_golden.py) carry the same pin to such machines.  CPU only."""
import ctypes

import numpy as np
import pytest

import bdb200 as bd
from cases import SMALL, build


def test_sfc64_and_normal_stream_bit_exact(oracle, reference):
    for seed in (0, 1, 123, 2**63 + 5):
        assert np.array_equal(oracle.sfc64(seed, 4096), reference.sfc64(seed, 4096))
    # 3M draws reach the wedge and tail branches of the ziggurat thousands of times
    assert np.array_equal(oracle.normals(123, 3_000_000), reference.normals(123, 3_000_000))


@pytest.mark.parametrize("kind,period", [(bd.BOX_OPEN, None), (bd.BOX_PERIODIC, (0.9, 1.0, 1.1)),
                                         (bd.BOX_PERIODIC, (0.5, 2.0, 0.31))])
@pytest.mark.parametrize("n", [0, 1, 100, 1000])
def test_buckets_and_pairs_bit_exact(oracle, reference, kind, period, n):
    x = np.random.default_rng(n + 7).normal(size=(n, 3)) * 1.3
    for spacing in (0.3, 0.45):
        assert np.array_equal(oracle.nsearch_buckets(kind, period, n, spacing, x),
                              reference.nsearch_buckets(kind, period, n, spacing, x))
        # same pairs in the same (bucket) order
        assert np.array_equal(oracle.nsearch_pairs(kind, period, n, spacing, x),
                              reference.nsearch_pairs(kind, period, n, spacing, x))


@pytest.mark.parametrize("name", sorted(SMALL))
def test_forces_energy_trajectory_bit_exact(oracle, reference, name):
    scen = SMALL[name]()
    so, ids = build(oracle, scen)
    sr, _ = build(reference, scen)
    assert np.array_equal(so.compute_forces(), sr.compute_forces())
    assert so.compute_energy() == sr.compute_energy()
    for cls, key in ((bd.FF_PAIR, "pair"), (bd.FF_BOND, "bond"), (bd.FF_BEND, "bend"),
                     (bd.FF_WALL, "wall"), (bd.FF_POINT, "point")):
        for ff in ids[key]:
            assert np.array_equal(so.compute_forces(0, cls, ff), sr.compute_forces(0, cls, ff)), (key, ff)
            assert so.compute_energy(0, cls, ff) == sr.compute_energy(0, cls, ff), (key, ff)
            if cls == bd.FF_WALL:
                assert np.array_equal(so.get_wall_stats(ff), sr.get_wall_stats(ff))
    assert np.array_equal(so.get_neighbor_list(), sr.get_neighbor_list())
    # 30 steps in two chunks (the second continues the RNG stream and the previous noise)
    for s in (so, sr):
        s.step(20, scen.timestep, scen.temperature, seed=5, first_step=1)
        s.step(10, scen.timestep, scen.temperature, seed=5, first_step=21)
    assert np.array_equal(so.get_positions(), sr.get_positions())


def test_replayed_noise_reproduces_reference_trajectory(oracle, reference):
    """The validation mode of the engine replays host-generated normals; check on the CPU that
    feeding the reference's own normal stream through `replay_normals` gives its trajectory."""
    scen = SMALL["ab_sphere"]()
    so, _ = build(oracle, scen)
    sr, _ = build(reference, scen)
    steps = 15
    z = reference.normals(77, steps * scen.n * 3)
    so.step(steps, scen.timestep, scen.temperature, seed=0, replay_normals=z)
    sr.step(steps, scen.timestep, scen.temperature, seed=77)
    assert np.array_equal(so.get_positions(), sr.get_positions())


def test_adaptive_timestep_bit_exact(oracle, reference):
    scen = SMALL["genome_10mb"]()
    so, _ = build(oracle, scen)
    sr, _ = build(reference, scen)
    for s in (so, sr):
        s.step(10, 1e-3, scen.temperature, seed=9, spacestep=0.01)
    assert np.array_equal(so.get_positions(), sr.get_positions())


def test_parameter_updates_follow_reference(oracle, reference):
    """bead_scale / bond_scale / wall semiaxes change every step in the 100kb driver
    (simulation_driver_interphase.cc:42-80)."""
    import scenarios as sc

    scen = SMALL["genome_100kb"]()
    so, ids = build(oracle, scen)
    sr, _ = build(reference, scen)
    semi = np.array(scen.params["semiaxes"])
    for k in range(5):
        t = sc.c4_tables(scen.params, bead_scale=0.8 + 0.02 * k, bond_scale=0.9 + 0.01 * k)
        semi = semi * 0.999
        for s in (so, sr):
            s.update_pair_ff(ids["pair"][0], t["pair"], t["dcut"])
            s.update_bond_ff(ids["bond"][0], t["chain_bond"])
            s.update_bond_ff(ids["bond"][1], t["loop_bond"])
            s.update_bond_ff(ids["bond"][2], t["nuc_bond"])
            s.update_wall_ff(ids["wall"][0], (0, 0, 0), semi, t["wall_in"])
            s.update_wall_ff(ids["wall"][1], (0, 0, 0), semi)
            s.step(1, scen.timestep, scen.temperature, seed=3, first_step=k + 1)
        assert np.array_equal(so.get_wall_stats(ids["wall"][0]), sr.get_wall_stats(ids["wall"][0]))
    assert np.array_equal(so.get_positions(), sr.get_positions())


def test_public_simulate_entry_point_matches_chunked_stepping(reference):
    """ref_step() drives md::detail::brownian_simulator directly; md::simulate_brownian_dynamics
    itself must give the same trajectory."""
    scen = SMALL["ab_box"]()
    a, _ = build(reference, scen)
    b, _ = build(reference, scen)
    a.step(12, scen.timestep, scen.temperature, seed=4)
    fn = reference.lib.ref_simulate
    fn.restype = ctypes.c_int
    fn.argtypes = [ctypes.c_void_p, ctypes.c_double, ctypes.c_double, ctypes.c_double, ctypes.c_int64, ctypes.c_uint64]
    assert fn(b.h, scen.temperature, scen.timestep, 0.0, 12, 4) == 0
    assert np.array_equal(a.get_positions(), b.get_positions())


def test_table_potential_equals_expression_templates(reference):
    fn = reference.lib.ref_selfcheck_expression_templates
    fn.restype = ctypes.c_double
    assert fn() == 0.0


def test_external_forces_are_one_more_forcefield(oracle, reference):
    """set_external_forces: a host-evaluated md::forcefield subclass (3-sim-1kb's glue / loop force
    fields) as one more member of the reference's sum forcefield."""
    scen = SMALL["chromatin_1kb"]()
    so, _ = build(oracle, scen)
    sr, _ = build(reference, scen)
    ext = np.random.default_rng(2).normal(size=(scen.n, 3)) * 5.0
    base = sr.compute_forces()
    for s in (so, sr):
        s.set_external_forces(ext)
    assert np.array_equal(sr.compute_forces(), base + ext)
    assert np.array_equal(so.compute_forces(), sr.compute_forces())
    # only the whole-system evaluation carries it
    assert np.array_equal(so.compute_forces(ff_class=1, ff=0), sr.compute_forces(ff_class=1, ff=0))
    z = reference.normals(3, 8 * scen.n * 3)
    so.step(8, scen.timestep, scen.temperature, replay_normals=z)
    sr.step(8, scen.timestep, scen.temperature, seed=3)
    assert np.array_equal(so.get_positions(), sr.get_positions())
    for s in (so, sr):
        s.set_external_forces(None)
    assert np.array_equal(so.compute_forces(), sr.compute_forces())


def test_dynamic_pair_lists_with_and_without_minimum_image(oracle, reference):
    """set_bond_pairs / set_bond_min_image: the oracle against micromd's bonded force field (plain
    displacement, the loop_forcefield case) and against a glue_forcefield-shaped class on
    md::periodic_box::shortest_displacement, with the list edited between steps."""
    import scenarios as sc
    from cases import rel_force_error

    rng = np.random.default_rng(17)
    s = sc.chromatin_1kb(n_chains=3, chain_len=120)
    L = np.asarray(s.period)
    n = s.n
    well = bd.pair_table(s.ntypes, lambda ti, tj: [bd.softwell(6, -1.5, 0.4)])
    spring = bd.pair_table(s.ntypes, lambda ti, tj: [bd.spring(40.0, 0.3)])

    def random_pairs(m):
        i = rng.integers(0, n, size=m)
        j = (i + rng.integers(1, n - 1, size=m)) % n
        return np.stack([i, j], axis=1).astype(np.uint32)

    systems = []
    for backend in (oracle, reference):
        e, _ = build(backend, s)
        systems.append(e)
    x = s.positions[0] + np.random.default_rng(5).integers(-2, 3, size=(n, 3)) * L   # images: min-image matters
    ids = []
    for e in systems:
        e.set_positions(x)
        glue = e.add_bond_ff(well, np.zeros((0, 2), np.uint32))
        loop = e.add_bond_ff(spring, np.zeros((0, 2), np.uint32))
        e.set_bond_min_image(glue, True)
        ids.append((glue, loop))
    for round_ in range(6):
        gp, lp = random_pairs(40 + 5 * round_), random_pairs(12)
        if round_ == 4:
            gp = gp[:0]                                   # an empty list is a valid list
        out = []
        for e, (glue, loop) in zip(systems, ids):
            e.set_bond_pairs(glue, gp)
            e.set_bond_pairs(loop, lp)
            fg = e.compute_forces(ff_class=bd.FF_BOND, ff=glue)
            fl = e.compute_forces(ff_class=bd.FF_BOND, ff=loop)
            eg = e.compute_energy(ff_class=bd.FF_BOND, ff=glue)
            e.step(4, s.timestep, s.temperature, seed=3, first_step=1 + 4 * round_)
            out.append((fg, fl, eg, e.get_positions()))
        (fg, fl, eg, xo), (rg, rl, reg, xr) = out
        assert rel_force_error(fl, rl) <= 1e-12
        if len(gp):
            assert np.abs(rg).max() > 0 and rel_force_error(fg, rg) <= 1e-12
        assert eg == pytest.approx(reg, rel=1e-12, abs=1e-12)
        assert np.abs(xo - xr).max() <= 1e-11
    for e in systems:
        e.close()
