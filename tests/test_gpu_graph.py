"""The CUDA-graph step path against the plain-launch path.  GPU only.

bd_step takes the graph path (one cudaGraphLaunch per step: k_rebuild_begin, the rebuild chain inside a
conditional node, the fused step kernel with its parameters patched per launch, k_wall_relax with the law and the
timestep baked in; bd_engine.cu `ensure_step_graph` / `launch_step_graph`) whenever the step is not adaptive, not
replayed, not kernel-timed and the bond rows are current -- i.e. in every production run -- while the oracle
trajectory tests replay normals and so never enter it.  Here the same Philox-driven run is made twice, with
`step_graph` = 0 and = 1, and everything observable must be IDENTICAL bit for bit: both paths launch the same
kernels on the same data in the same order, so any difference is a wrong graph (a missing node in the conditional
body, a stale baked argument, a signature that failed to notice a change).

Covered: an open box with per-replica wall relaxation and a periodic box; stretches long enough for rebuilds
inside the conditional node; a new bond list between calls (the step after it leaves the graph path once, then
returns to it); a different timestep (baked into the relaxation node); a different list radius (new row arrays:
new signature); a forcefield switched off and on again (launch parameters only).
"""
import numpy as np
import pytest

import bdb200 as bd
import scenarios as sc

pytestmark = pytest.mark.gpu

SPRING = np.array([1.0e4, 1.2e4, 0.9e4])
MOBILITY = 1.0e-4


def observe(system, ids, replicas):
    out = {"x": np.stack([system.get_positions(r) for r in range(replicas)])}
    c = system.get_counters()
    out["counters"] = np.array([c.steps_done, c.list_rebuilds, c.overflow_beads])
    for w in ids["wall"]:
        out[f"wall{w}.stats"] = system.get_wall_stats_all(w)
        out[f"wall{w}.geom"] = np.stack([np.concatenate(system.get_wall_geometry(w, r)) for r in range(replicas)])
    return out


def assert_identical(a, b, where):
    assert a.keys() == b.keys()
    for k in a:
        assert np.array_equal(a[k], b[k]), (where, k, np.abs(np.asarray(a[k], float) - np.asarray(b[k], float)).max())


def run_script(engine, scen, graph, script):
    """script(system, ids, checkpoint): drives the system; checkpoint(label) records everything observable"""
    system = bd.System(engine, scen.n, scen.replicas)
    ids = scen.apply(system)
    system.set_option("step_graph", graph)
    assert system.get_option("step_graph") == graph
    marks = []
    script(system, ids, lambda label: marks.append((label, observe(system, ids, scen.replicas))))
    return marks


def both_ways(engine, scen, script):
    plain = run_script(engine, scen, 0, script)
    graphed = run_script(engine, scen, 1, script)
    assert [m[0] for m in plain] == [m[0] for m in graphed]
    for (label, a), (_, b) in zip(plain, graphed):
        assert_identical(a, b, label)
    return plain


def test_open_box_with_wall_relaxation(engine):
    R = 3
    scen = sc.genome_100kb(3000, 30, seed=9, replicas=R, bead_scale=0.9, bond_scale=0.95)
    semi0 = np.array(scen.params["semiaxes"])
    dt, kT = scen.timestep, scen.temperature

    def script(s, ids, checkpoint):
        for r in range(R):  # every replica its own squeezed ellipsoid, so that the walls are pressed on and move
            for w in ids["wall"]:
                s.update_wall_ff(w, (0, 0, 0), semi0 * (0.86 + 0.02 * r), replica=r)
        s.set_wall_relaxation(ids["wall"], SPRING, MOBILITY)
        s.step(150, dt, kT, seed=21)
        checkpoint("first stretch")
        s.step(150, dt, kT, seed=21, first_step=151)
        checkpoint("second stretch")
        # a new bond list: the next step rebuilds the bond rows outside the graph, later steps are graphed again
        pairs = scen.bond_ffs[0]["pairs"]
        s.set_bond_pairs(ids["bond"][0], pairs[: len(pairs) - 40])
        s.step(60, dt, kT, seed=21, first_step=301)
        checkpoint("after set_bond_pairs")
        # another timestep: baked into the relaxation node of the graph
        s.step(60, 0.5 * dt, kT, seed=21, first_step=361)
        checkpoint("after a timestep change")
        # another list radius: new rows, the graph's signature must notice
        s.set_option("verlet_factor", 1.3)
        s.step(60, dt, kT, seed=21, first_step=421)
        checkpoint("after verlet_factor")
        s.set_ff_enabled(bd.FF_PAIR, ids["pair"][0], False)
        s.step(20, dt, kT, seed=21, first_step=481)
        s.set_ff_enabled(bd.FF_PAIR, ids["pair"][0], True)
        s.step(20, dt, kT, seed=21, first_step=501)
        checkpoint("after a forcefield was off and on")
        s.set_wall_relaxation([], SPRING, MOBILITY)  # relaxation off: a graph without the relax node
        s.step(30, dt, kT, seed=21, first_step=521)
        checkpoint("after the relaxation was switched off")

    marks = both_ways(engine, scen, script)
    first, last = marks[0][1], marks[-1][1]
    assert first["counters"][1] >= 2, "no list rebuild happened inside the first stretch"
    assert marks[1][1]["counters"][1] > first["counters"][1], "no rebuild inside the second stretch"
    w = [k for k in first if k.endswith(".geom")][0]
    assert np.abs(first[w][:, 3:] - semi0 * 0.86).max() > 1e-9  # the walls moved
    assert np.isfinite(last["x"]).all()


def test_periodic_box(engine):
    scen = sc.ab_box(60, 20, seed=4, replicas=2)
    dt, kT = scen.timestep, scen.temperature

    def script(s, ids, checkpoint):
        s.step(200, dt, kT, seed=5)
        checkpoint("first stretch")
        s.step(200, dt, kT, seed=5, first_step=201)
        checkpoint("second stretch")
        if ids["bond"]:
            pairs = scen.bond_ffs[0]["pairs"]
            s.set_bond_pairs(ids["bond"][0], pairs[: len(pairs) - 10])
        s.step(100, 0.7 * dt, kT, seed=6, first_step=401)
        checkpoint("after set_bond_pairs, another timestep and seed")

    marks = both_ways(engine, scen, script)
    assert marks[1][1]["counters"][1] >= 2, "no list rebuild inside the run"


def test_graph_path_is_the_one_taken(engine):
    """The production default is the graph path: fewer host launches than steps x kernels, and step_graph = 0 turns it off."""
    scen = sc.ab_box(60, 20, seed=4)
    launches = {}
    for graph in (1, 0):
        s = bd.System(engine, scen.n)
        scen.apply(s)
        default = s.get_option("step_graph")
        assert default == 1
        s.set_option("step_graph", graph)
        s.step(5, scen.timestep, scen.temperature, seed=1)  # list build + warm-up
        before = s.get_launch_count()
        s.step(40, scen.timestep, scen.temperature, seed=1, first_step=6)
        launches[graph] = s.get_launch_count() - before
    assert launches[1] > 0 and launches[0] > 0 and launches[1] != launches[0]
