#!/usr/bin/env python
"""Full-size fixtures for the configurations BASELINE.json names, from the UNMODIFIED reference
(oracle/_ref/libmicromd_ref.so).  Run where /root/reference exists:

    python tests/golden/make_named_golden.py

A 62 600-bead force array is 1.5 MB and a million-bead one 24 MB, so tests/golden/named_sizes.npz keeps,
per configuration: the sha256 of the reference's full arrays (the oracle reproduces the reference bit for
bit, tests/test_oracle_golden.py, so it must reproduce the hashes), the values at 1 000 sampled beads
(what the engine is held to, with a tolerance), the total energy, and count / sum / xor of the 64-bit keys
of the Verlet pair set (bit-exact for everybody).
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
for p in ("2022a-genome-dynamics_b200", "tests", "oracle"):
    sys.path.insert(0, os.path.join(ROOT, p))

import checkers  # noqa: E402
from cases import NAMED, NAMED_SEED, NAMED_STEPS, build, digest, pair_digest, sample_beads  # noqa: E402


def main():
    ref = checkers.load_reference()
    out = {}
    for name, make in NAMED.items():
        scen = make()
        s, _ = build(ref, scen)
        idx = sample_beads(scen.n)
        f = s.compute_forces()
        out[f"{name}.n"] = scen.n
        out[f"{name}.positions_sha"] = digest(scen.positions[0])
        out[f"{name}.forces_sha"] = digest(f)
        out[f"{name}.forces_sample"] = f[idx]
        out[f"{name}.force_scale"] = np.linalg.norm(f, axis=1).max()
        out[f"{name}.energy"] = s.compute_energy()
        out[f"{name}.pairs"] = pair_digest(s.get_neighbor_list())
        s.step(NAMED_STEPS, scen.timestep, scen.temperature, seed=NAMED_SEED)
        x = s.get_positions()
        out[f"{name}.after_sha"] = digest(x)
        out[f"{name}.after_sample"] = x[idx]
        print(name, scen.n, "beads", int(out[f"{name}.pairs"][0]), "listed pairs", flush=True)
        s.close()
    np.savez_compressed(os.path.join(HERE, "named_sizes.npz"), **out)


if __name__ == "__main__":
    main()
