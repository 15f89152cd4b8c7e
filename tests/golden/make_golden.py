#!/usr/bin/env python
"""Generates tests/golden/*.npz from the UNMODIFIED reference (oracle/_ref/libmicromd_ref.so, built
by `make -C oracle` from /root/reference).  Run here (where /root/reference exists):

    python tests/golden/make_golden.py

The fixtures pin the oracle and the engine on machines without the reference tree (the GPU box).
Every array is produced by reference code only; inputs come from scenarios.py with fixed seeds.
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, os.path.join(ROOT, "2022a-genome-dynamics_b200"))
sys.path.insert(0, os.path.join(ROOT, "tests"))
sys.path.insert(0, os.path.join(ROOT, "oracle"))

import bdb200 as bd  # noqa: E402
import checkers  # noqa: E402
from cases import GOLDEN, GOLDEN_SEED, GOLDEN_STEPS, build  # noqa: E402


def main():
    ref = checkers.load_reference()

    # RNG stream: first 4096 sfc64 outputs and normals, plus a checksum over 1M normals
    z = ref.normals(GOLDEN_SEED, 1_000_000)
    np.savez_compressed(os.path.join(HERE, "rng.npz"), seed=GOLDEN_SEED,
                        sfc64=ref.sfc64(GOLDEN_SEED, 4096), normals=z[:4096],
                        normals_sum=z.sum(), normals_sumsq=(z * z).sum(),
                        normals_xor=np.bitwise_xor.reduce(z.view(np.uint64)))

    # neighbour search: the reference's own unit-test shapes (100 points, dcut 0.3) + a larger set
    out = {}
    for tag, n, scale in (("t100", 100, 1.0), ("t1000", 1000, 1.3)):
        x = np.random.default_rng(n).normal(size=(n, 3)) * scale
        out[f"{tag}_x"] = x
        for kname, kind, period in (("open", bd.BOX_OPEN, None), ("per", bd.BOX_PERIODIC, (0.9, 1.0, 1.1))):
            out[f"{tag}_{kname}_buckets"] = ref.nsearch_buckets(kind, period, n, 0.3, x)
            out[f"{tag}_{kname}_pairs"] = ref.nsearch_pairs(kind, period, n, 0.3, x)
    np.savez_compressed(os.path.join(HERE, "nsearch.npz"), **out)

    # per-scenario forces, energies, wall statistics, Verlet list and a short trajectory
    for name, make in GOLDEN.items():
        scen = make()
        s, ids = build(ref, scen)
        rec = dict(positions0=scen.positions[0], forces=s.compute_forces(), energy=s.compute_energy(),
                   neighbor_list=s.get_neighbor_list())
        for cls, key in ((bd.FF_PAIR, "pair"), (bd.FF_BOND, "bond"), (bd.FF_BEND, "bend"),
                         (bd.FF_WALL, "wall"), (bd.FF_POINT, "point")):
            for ff in ids[key]:
                rec[f"forces_{key}{ff}"] = s.compute_forces(0, cls, ff)
                rec[f"energy_{key}{ff}"] = s.compute_energy(0, cls, ff)
                if cls == bd.FF_WALL:
                    rec[f"stats_{key}{ff}"] = s.get_wall_stats(ff)
        s.step(GOLDEN_STEPS, scen.timestep, scen.temperature, seed=GOLDEN_SEED)
        rec["positions_after"] = s.get_positions()
        rec["normals"] = ref.normals(GOLDEN_SEED, GOLDEN_STEPS * scen.n * 3)
        rec["energy_after"] = s.compute_energy()
        np.savez_compressed(os.path.join(HERE, f"{name}.npz"), **rec)
        print(name, scen.n, "beads,", len(rec["neighbor_list"]), "listed pairs")


if __name__ == "__main__":
    main()
