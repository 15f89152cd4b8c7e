"""Synthetic systems shaped like the reference drivers' setups (BASELINE.md
section 3).  The real genome tables and Hi-C blobs are not in the reference tree, so every input
here is generated from a fixed seed; the same arrays are handed to the engine, the oracle and the
reference harness so all three see bit-identical fp64 coordinates and parameters.

Each builder returns a `Scenario`; `Scenario.apply(system)` performs, through the C ABI, the calls
the corresponding driver makes through micromd:

* C1 ab_box      4-sim-ab/box/src/simulation/simulation_driver.cc:93-141
* C2 ab_sphere   4-sim-ab/sphere/src/simulation_driver.cc:95-228
* C3 genome_10mb 5-sim-genome/src/simulation_spindle/simulation_driver.cc:90-172
* C4 genome_100kb 5-sim-genome/src/simulation_interphase/simulation_driver_forcefield.cc:19-235
* C5 chromatin_1kb 3-sim-1kb/src/simulation/simulation.cpp:100-184
"""
from __future__ import annotations

from dataclasses import dataclass, field
from typing import Dict, List, Optional, Tuple

import numpy as np

import bdb200 as bd


@dataclass
class Scenario:
    name: str
    n: int
    types: np.ndarray                 # uint8 [n]
    mobility: np.ndarray              # [ntypes]
    positions: np.ndarray             # [replicas, n, 3] fp64
    box_kind: int = bd.BOX_OPEN
    period: Tuple[float, float, float] = (1.0, 1.0, 1.0)
    pair_ffs: List[dict] = field(default_factory=list)    # {table, dcut, targets}
    bond_ffs: List[dict] = field(default_factory=list)    # {table, pairs}
    bend_ffs: List[dict] = field(default_factory=list)    # {triples, energies}
    wall_ffs: List[dict] = field(default_factory=list)    # {shape, direction, table, center, semiaxes}
    point_ffs: List[dict] = field(default_factory=list)   # {table, source, targets}
    temperature: float = 1.0
    timestep: float = 1e-5
    chains: Optional[np.ndarray] = None                    # [nchains, 2] (start, end)
    params: Dict[str, float] = field(default_factory=dict)

    @property
    def ntypes(self) -> int:
        return len(self.mobility)

    @property
    def replicas(self) -> int:
        return self.positions.shape[0]

    def apply(self, system: bd.System, replicas: Optional[range] = None) -> Dict[str, List[int]]:
        """Configure `system` (any backend).  Returns the forcefield ids by class."""
        system.set_box(self.box_kind, self.period)
        system.set_types(self.types, self.mobility)
        ids: Dict[str, List[int]] = {"pair": [], "bond": [], "bend": [], "wall": [], "point": []}
        for ff in self.pair_ffs:
            ids["pair"].append(system.add_pair_ff(ff["table"], ff["dcut"], ff.get("targets")))
        for ff in self.bond_ffs:
            ids["bond"].append(system.add_bond_ff(ff["table"], ff["pairs"]))
        for ff in self.bend_ffs:
            ids["bend"].append(system.add_bend_ff(ff["triples"], ff["energies"]))
        for ff in self.wall_ffs:
            ids["wall"].append(
                system.add_wall_ff(ff["shape"], ff["direction"], ff["table"], ff["center"], ff["semiaxes"])
            )
        for ff in self.point_ffs:
            ids["point"].append(system.add_point_ff(ff["table"], ff["source"], ff.get("targets")))
        reps = replicas if replicas is not None else range(min(system.replicas, self.replicas))
        for k, r in enumerate(reps):
            system.set_positions(self.positions[r], replica=k)
        return ids


# ---- geometry helpers ---------------------------------------------------------------------------


def chain_ranges(lengths) -> np.ndarray:
    ends = np.cumsum(lengths)
    starts = ends - np.asarray(lengths)
    return np.stack([starts, ends], axis=1).astype(np.int64)


def range_bonds(chains: np.ndarray, stride: int = 1) -> np.ndarray:
    """add_bonded_range (stride 1) or the (i, i+2) loop bonds (stride 2) of every chain."""
    out = []
    for s, e in chains:
        i = np.arange(s, max(s, e - stride))
        out.append(np.stack([i, i + stride], axis=1))
    return np.concatenate(out).astype(np.uint32) if out else np.zeros((0, 2), np.uint32)


def range_triples(chains: np.ndarray) -> np.ndarray:
    out = []
    for s, e in chains:
        i = np.arange(s, max(s, e - 2))
        out.append(np.stack([i, i + 1, i + 2], axis=1))
    return np.concatenate(out).astype(np.uint32) if out else np.zeros((0, 3), np.uint32)


def random_walk_chains(rng, chains, step, replicas, confine_radius=None, box=None, start_spread=None):
    """Random-walk chains; optionally reflected into a sphere of `confine_radius`.

    Vectorised over chains and replicas, looping over the position along the chain.
    """
    n = int(chains[-1, 1])
    nch = len(chains)
    maxlen = int((chains[:, 1] - chains[:, 0]).max())
    pos = np.zeros((replicas, n, 3))
    if confine_radius is not None:
        spread = start_spread if start_spread is not None else 0.6 * confine_radius
        start = rng.normal(size=(replicas, nch, 3))
        start *= (spread * rng.random((replicas, nch, 1)) ** (1 / 3)) / np.linalg.norm(start, axis=2, keepdims=True)
    else:
        L = np.asarray(box)
        start = rng.random((replicas, nch, 3)) * L
    cur = start
    lengths = chains[:, 1] - chains[:, 0]
    for k in range(maxlen):
        alive = lengths > k
        idx = chains[alive, 0] + k
        pos[:, idx, :] = cur[:, alive, :]
        d = rng.normal(size=(replicas, nch, 3))
        d *= step / np.linalg.norm(d, axis=2, keepdims=True)
        nxt = cur + d
        if confine_radius is not None:
            r = np.linalg.norm(nxt, axis=2, keepdims=True)
            out = r > confine_radius * 0.98
            nxt = np.where(out, cur - d, nxt)  # bounce back instead of leaving the sphere
        cur = nxt
    return pos


# ---- C1 / C2: 4-sim-ab --------------------------------------------------------------------------


def _ab_common(n_chains, chain_len, seed, replicas, confined):
    rng = np.random.default_rng(seed)
    n = n_chains * chain_len
    p = dict(a_core_diameter=0.30, b_core_diameter=0.24, a_core_repulsion=2.0, b_core_repulsion=2.0,
             bond_spring=70.0, mobility=1.0)
    chains = chain_ranges([chain_len] * n_chains)
    # alternating pure-A / pure-B chains: (a, b) in {(1,0), (0,1)}
    types = np.repeat(np.arange(n_chains) % 2, chain_len).astype(np.uint8)
    ab = np.array([[1.0, 0.0], [0.0, 1.0]])
    volume = n * np.pi / 6 * p["a_core_diameter"] ** 3 / 0.2  # volume fraction 0.2

    def pair(ti, tj):
        a = (ab[ti, 0] + ab[tj, 0]) / 2
        b = (ab[ti, 1] + ab[tj, 1]) / 2
        return [bd.softcore(2, 3, p["a_core_repulsion"], p["a_core_diameter"], factor=a),
                bd.softcore(8, 3, p["b_core_repulsion"], p["b_core_diameter"], factor=b)]

    pair_ff = dict(table=bd.pair_table(2, pair), dcut=max(p["a_core_diameter"], p["b_core_diameter"]))
    bond_ff = dict(table=bd.pair_table(2, lambda ti, tj: [bd.harmonic(p["bond_spring"])]),
                   pairs=range_bonds(chains))
    return rng, n, p, chains, types, ab, volume, pair_ff, bond_ff


def ab_box(n_chains=200, chain_len=20, seed=1, replicas=1) -> Scenario:
    rng, n, p, chains, types, ab, volume, pair_ff, bond_ff = _ab_common(n_chains, chain_len, seed, replicas, False)
    L = volume ** (1 / 3)
    pos = random_walk_chains(rng, chains, 0.2, replicas, box=(L, L, L))
    return Scenario("ab_box", n, types, np.array([1.0, 1.0]), pos, bd.BOX_PERIODIC, (L, L, L),
                    pair_ffs=[pair_ff], bond_ffs=[bond_ff], chains=chains, params=p)


def ab_sphere(n_chains=200, chain_len=20, seed=2, replicas=1, inner_wall=False) -> Scenario:
    rng, n, p, chains, types, ab, volume, pair_ff, bond_ff = _ab_common(n_chains, chain_len, seed, replicas, True)
    R = (volume * 3 / (4 * np.pi)) ** (1 / 3)
    pos = random_walk_chains(rng, chains, 0.2, replicas, confine_radius=R)
    wall_a, wall_b, mult, spring = 0.0, 1.0, 1.0, 1.0

    def inward(ti):
        a = (ab[ti, 0] + wall_a) / 2
        b = (ab[ti, 1] + wall_b) / 2
        return [bd.softcore(2, 3, p["a_core_repulsion"], p["a_core_diameter"] / 2, factor=mult * a),
                bd.softcore(8, 3, p["b_core_repulsion"], p["b_core_diameter"] / 2, factor=mult * b)]

    walls = [
        dict(shape=bd.WALL_SPHERE, direction=bd.WALL_INWARD, table=bd.field_table(2, inward),
             center=(0, 0, 0), semiaxes=(R, R, R)),
        dict(shape=bd.WALL_SPHERE, direction=bd.WALL_OUTWARD,
             table=bd.field_table(2, lambda t: [bd.harmonic(spring)]), center=(0, 0, 0), semiaxes=(R, R, R)),
    ]
    if inner_wall:
        Ri = 0.3 * R
        walls += [
            dict(shape=bd.WALL_SPHERE, direction=bd.WALL_INWARD,
                 table=bd.field_table(2, lambda t: [bd.harmonic(spring)]), center=(0, 0, 0), semiaxes=(Ri, Ri, Ri)),
            dict(shape=bd.WALL_SPHERE, direction=bd.WALL_OUTWARD, table=bd.field_table(2, inward),
                 center=(0, 0, 0), semiaxes=(Ri, Ri, Ri)),
        ]
    return Scenario("ab_sphere", n, types, np.array([1.0, 1.0]), pos, bd.BOX_OPEN,
                    pair_ffs=[pair_ff], bond_ffs=[bond_ff], wall_ffs=walls, chains=chains, params=p)


# ---- C3: 10Mb ana/telophase ---------------------------------------------------------------------

# hg19 chromosome lengths in Mb (1..22, X), used only as proportions for synthetic chain lengths
_HG19_MB = [249, 243, 198, 191, 181, 171, 159, 146, 141, 136, 135, 134, 115, 107, 103, 90, 81, 78,
            59, 63, 48, 51, 155]


def genome_chain_lengths(total: int) -> List[int]:
    w = np.array(_HG19_MB + _HG19_MB, dtype=np.float64)
    lengths = np.maximum(4, np.floor(w / w.sum() * total)).astype(int)
    lengths[0] += total - lengths.sum()
    return lengths.tolist()


def genome_10mb(total=640, seed=3, replicas=1, bend_energy=0.0, packing=True) -> Scenario:
    rng = np.random.default_rng(seed)
    chains = chain_ranges(genome_chain_lengths(total))
    n = int(chains[-1, 1])
    p = dict(init_bead_diameter=0.2, init_bead_repulsion=5.0, init_bond_length=0.2, init_bond_spring=500.0,
             init_spindle_spring=1.0, init_packing_radius=1.0, init_packing_spring=1.0)
    # rods in random directions from a blob around (5,0,0), as run_initialization() does
    pos = np.zeros((replicas, n, 3))
    for r in range(replicas):
        for s, e in chains:
            centroid = np.array([5.0, 0, 0]) + rng.normal(size=3)
            step = rng.normal(size=3)
            step *= p["init_bond_length"] / np.linalg.norm(step)
            k = np.arange(e - s)[:, None]
            pos[r, s:e] = centroid - step * (e - s) / 2 + k * step
    cent = np.zeros(n, dtype=np.uint8)
    for s, e in chains:
        c = (s + e) // 2
        cent[c - 1:c + 2] = 1
    one = lambda terms: bd.table([terms])
    sc = Scenario(
        "genome_10mb", n, np.zeros(n, np.uint8), np.array([1.0]), pos, bd.BOX_OPEN,
        pair_ffs=[dict(table=one([bd.softcore(2, 3, p["init_bead_repulsion"], p["init_bead_diameter"])]),
                       dcut=p["init_bead_diameter"])],
        bond_ffs=[dict(table=one([bd.semispring(p["init_bond_spring"], p["init_bond_length"])]),
                       pairs=range_bonds(chains))],
        point_ffs=[dict(table=one([bd.harmonic(p["init_spindle_spring"])]), source=(0, 0, 0), targets=cent)],
        temperature=0.1, timestep=1e-4, chains=chains, params=p)
    if bend_energy != 0.0:
        tr = range_triples(chains)
        sc.bend_ffs.append(dict(triples=tr, energies=np.full(len(tr), bend_energy)))
    if packing:
        sc.point_ffs.append(dict(table=one([bd.semispring(p["init_packing_spring"], p["init_packing_radius"])]),
                                 source=(0, 0, 0), targets=None))
    return sc


# ---- C4: 100kb interphase -----------------------------------------------------------------------

C4_PARAMS = dict(
    a_core_diameter=0.30, b_core_diameter=0.24, a_core_repulsion=2.0, b_core_repulsion=2.0,
    a_core_bond_spring=60.0, a_core_bond_length=0.30, b_core_bond_spring=90.0, b_core_bond_length=0.24,
    a_core_2nd_bond_spring=4.0, b_core_2nd_bond_spring=8.0,
    nucleolus_a_factor=5.0, nucleolus_b_factor=5.0, nucleolus_bond_spring=5.0, nucleolus_bond_length=0.0,
    wall_packing_spring=5000.0, wall_a_factor=5.0, wall_b_factor=5.0, wall_mobility=1e-4,
    wall_semiaxes_spring=1e4,
)
# bead classes: A, B, u (0.5, 0.5), nucleolar side-bead (5, 5)
C4_AB = np.array([[1.0, 0.0], [0.0, 1.0], [0.5, 0.5], [5.0, 5.0]])


def c4_tables(p, bead_scale=1.0, bond_scale=1.0):
    """Type tables of the five interphase forcefields for the current (bead_scale, bond_scale);
    the per-step callback changes both (simulation_driver_interphase.cc:59-67)."""
    ab = C4_AB

    def pair(ti, tj):
        a = 0.5 * (ab[ti, 0] + ab[tj, 0])
        b = 0.5 * (ab[ti, 1] + ab[tj, 1])
        return [bd.softcore(2, 3, p["a_core_repulsion"], p["a_core_diameter"] * bead_scale, factor=a),
                bd.softcore(8, 3, p["b_core_repulsion"], p["b_core_diameter"] * bead_scale, factor=b)]

    def chain_bond(ti, tj):
        a = 0.5 * (ab[ti, 0] + ab[tj, 0])
        b = 0.5 * (ab[ti, 1] + ab[tj, 1])
        K = a * p["a_core_bond_spring"] + b * p["b_core_bond_spring"]
        l = a * p["a_core_bond_length"] + b * p["b_core_bond_length"]
        s = bond_scale
        return [bd.semispring(K * (1 / (s * s)), l * s)]

    def loop_bond(ti, tj):
        a = 0.5 * (ab[ti, 0] + ab[tj, 0])
        b = 0.5 * (ab[ti, 1] + ab[tj, 1])
        K = a * p["a_core_2nd_bond_spring"] + b * p["b_core_2nd_bond_spring"]
        s = bond_scale
        return [bd.harmonic(K * (1 / (s * s)))]

    def nuc_bond(ti, tj):
        s = bond_scale
        return [bd.semispring(p["nucleolus_bond_spring"] * (1 / (s * s)), p["nucleolus_bond_length"] * s)]

    def wall_in(ti):
        a = 0.5 * (ab[ti, 0] + p["wall_a_factor"])
        b = 0.5 * (ab[ti, 1] + p["wall_b_factor"])
        return [bd.softcore(2, 3, p["a_core_repulsion"], p["a_core_diameter"] / 2 * bead_scale, factor=a),
                bd.softcore(8, 3, p["b_core_repulsion"], p["b_core_diameter"] / 2 * bead_scale, factor=b)]

    T = len(ab)
    return dict(
        pair=bd.pair_table(T, pair),
        dcut=max(p["a_core_diameter"], p["b_core_diameter"]) * bead_scale,
        chain_bond=bd.pair_table(T, chain_bond),
        loop_bond=bd.pair_table(T, loop_bond),
        nuc_bond=bd.pair_table(T, nuc_bond),
        wall_in=bd.field_table(T, wall_in),
        wall_out=bd.field_table(T, lambda t: [bd.harmonic(p["wall_packing_spring"])]),
    )


def genome_100kb(total=62000, n_nor=300, seed=4, replicas=1, bead_scale=1.0, bond_scale=1.0,
                 volume_fraction=0.2) -> Scenario:
    rng = np.random.default_rng(seed)
    p = dict(C4_PARAMS)
    chains = chain_ranges(genome_chain_lengths(total))
    n_chrom = int(chains[-1, 1])
    n_nor = min(n_nor, n_chrom // 8)
    n = n_chrom + 2 * n_nor
    # A/B/u in ratio 4:4:2, in runs of geometric mean length 10
    types = np.zeros(n, dtype=np.uint8)
    i = 0
    while i < n_chrom:
        run = int(rng.geometric(0.1))
        types[i:i + run] = rng.choice(3, p=[0.4, 0.4, 0.2])
        i += run
    types[n_chrom:] = 3
    nor = np.sort(rng.choice(n_chrom, size=n_nor, replace=False))
    side = n_chrom + np.arange(2 * n_nor)
    nuc_pairs = np.stack([np.repeat(nor, 2), side], axis=1).astype(np.uint32)

    volume = n * np.pi / 6 * p["a_core_diameter"] ** 3 / volume_fraction
    R = (volume * 3 / (4 * np.pi)) ** (1 / 3)
    semiaxes = (1.15 * R, 1.0 * R, R / 1.15)
    Rmin = min(semiaxes)
    pos = np.zeros((replicas, n, 3))
    pos[:, :n_chrom] = random_walk_chains(rng, chains, 0.25, replicas, confine_radius=Rmin)
    pos[:, side] = pos[:, np.repeat(nor, 2)] + 0.05 * rng.normal(size=(replicas, 2 * n_nor, 3))

    t = c4_tables(p, bead_scale, bond_scale)
    sc = Scenario(
        "genome_100kb", n, types, np.ones(len(C4_AB)), pos, bd.BOX_OPEN,
        pair_ffs=[dict(table=t["pair"], dcut=t["dcut"])],
        bond_ffs=[dict(table=t["chain_bond"], pairs=range_bonds(chains)),
                  dict(table=t["loop_bond"], pairs=range_bonds(chains, 2)),
                  dict(table=t["nuc_bond"], pairs=nuc_pairs)],
        wall_ffs=[dict(shape=bd.WALL_ELLIPSOID, direction=bd.WALL_INWARD, table=t["wall_in"],
                       center=(0, 0, 0), semiaxes=semiaxes),
                  dict(shape=bd.WALL_ELLIPSOID, direction=bd.WALL_OUTWARD, table=t["wall_out"],
                       center=(0, 0, 0), semiaxes=semiaxes)],
        temperature=1.0, timestep=1e-5, chains=chains, params=p)
    sc.params["semiaxes"] = semiaxes
    return sc


# ---- C5: 1kb chromatin --------------------------------------------------------------------------


def chromatin_1kb(n_chains=6, chain_len=1000, seed=5, replicas=1, volume_fraction=0.1) -> Scenario:
    rng = np.random.default_rng(seed)
    n = n_chains * chain_len
    p = dict(repulsive_energy=3.0, repulsive_diameter=0.30, attractive_energy=0.3, attractive_diameter=0.45,
             bond_spring=100.0, bond_length=0.30, bending_energy=1.0, glue_energy=2.0, glue_distance=0.45,
             loop_spring=50.0)
    chains = chain_ranges([chain_len] * n_chains)
    volume = n * np.pi / 6 * p["repulsive_diameter"] ** 3 / volume_fraction
    L = volume ** (1 / 3)
    pos = random_walk_chains(rng, chains, p["bond_length"], replicas, box=(L, L, L))
    one = lambda terms: bd.table([terms])
    tr = range_triples(chains)
    # static "glue" (HP1 bridges) and cohesin loop pair sets, as the callback would maintain them
    n_glue, n_loop = n // 20, n // 200
    gi = rng.choice(n - 1, size=n_glue, replace=False)
    glue_pairs = np.stack([gi, np.minimum(n - 1, gi + rng.integers(3, 30, size=n_glue))], axis=1).astype(np.uint32)
    # loops are extruded along one chain (3-sim-1kb/src/simulation/loops/basic_loop_simulator.cpp:165-316):
    # both anchors stay inside the chain of the first one
    li = rng.choice(n - 1, size=max(1, n_loop), replace=False)
    chain_end = chains[np.searchsorted(chains[:, 1], li, side="right"), 1]
    lj = np.minimum(li + rng.integers(20, 200, size=len(li)), chain_end - 1)
    keep = lj > li
    loop_pairs = np.stack([li[keep], lj[keep]], axis=1).astype(np.uint32)
    return Scenario(
        "chromatin_1kb", n, np.zeros(n, np.uint8), np.array([1.0]), pos, bd.BOX_PERIODIC, (L, L, L),
        pair_ffs=[dict(table=one([bd.softcore(2, 3, p["repulsive_energy"], p["repulsive_diameter"]),
                                  bd.softcore(8, 3, -p["attractive_energy"], p["attractive_diameter"])]),
                       dcut=max(p["repulsive_diameter"], p["attractive_diameter"]))],
        bond_ffs=[dict(table=one([bd.spring(p["bond_spring"], p["bond_length"])]), pairs=range_bonds(chains)),
                  dict(table=one([bd.softcore(8, 3, -p["glue_energy"], p["glue_distance"])]), pairs=glue_pairs),
                  dict(table=one([bd.spring(p["loop_spring"], p["repulsive_diameter"])]), pairs=loop_pairs)],
        bend_ffs=[dict(triples=tr, energies=np.full(len(tr), p["bending_energy"]))],
        temperature=1.0, timestep=1e-5, chains=chains, params=p)


def random_points(n, seed=0, box_kind=bd.BOX_OPEN, period=(0.9, 1.0, 1.1)):
    """The neighbour-search unit-test inputs of the reference
    (micromd tests/misc/test_neighbor_searcher.cc:54-127): N(0,1) points, dcut 0.3."""
    rng = np.random.default_rng(seed)
    return rng.normal(size=(n, 3))
