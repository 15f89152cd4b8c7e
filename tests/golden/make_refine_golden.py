#!/usr/bin/env python
"""Golden vectors for the spline refinement, made by the REFERENCE's own module.

Imports 5-sim-genome/src/refine/refinement.py from /root/reference (it needs only numpy + scipy; run.py's
h5py store is not involved) and restates run.py:27-40's loop over chromosomes on synthetic coarse paths.
Run here (the reference tree does not exist on the GPU box); the output is committed.

    python tests/golden/make_refine_golden.py
"""
import importlib.util
import os

import numpy as np

REF = "/root/reference/5-sim-genome/src/refine/refinement.py"
OUT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "refine_spline.npz")


def main():
    spec = importlib.util.spec_from_file_location("reference_refinement", REF)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    do_refine = mod.get_refinement_method("spline")

    rng = np.random.default_rng(20221)
    factor = 100                                    # init_coarse_graining (config_entries.inc:46)
    lengths = [4, 5, 6, 9, 14, 25, 16, 7]           # coarse beads per chain; 4 is splprep's minimum (m > k)
    coarse_ranges, fine_ranges, coarse = [], [], []
    beg = fbeg = 0
    for m in lengths:
        # a confined random walk, like the packed 10Mb conformation
        steps = rng.normal(size=(m, 3)) * 0.3
        coarse.append(np.cumsum(steps, axis=0) + rng.normal(size=3))
        coarse_ranges.append((beg, beg + m))
        # the fine chain is up to one coarse bead shorter than m * factor (run.py:34 truncates)
        nf = m * factor - int(rng.integers(0, factor))
        fine_ranges.append((fbeg, fbeg + nf))
        beg += m
        fbeg += nf
    coarse = np.concatenate(coarse)
    n_chrom = fbeg
    attach = np.stack([rng.choice(n_chrom, size=12, replace=False), n_chrom + np.arange(12)], axis=1)
    n_fine = n_chrom + 12

    fine = np.empty((n_fine, 3))
    for (b, e), (fb, fe) in zip(coarse_ranges, fine_ranges):
        chain = do_refine(coarse[b:e], (e - b) * factor)
        fine[fb:fe] = chain[: fe - fb]
    for nor, nuc in attach:
        fine[nuc] = fine[nor]
    np.savez_compressed(OUT, coarse=coarse, coarse_ranges=np.array(coarse_ranges, dtype=np.int64),
                        fine_ranges=np.array(fine_ranges, dtype=np.int64), attach=attach.astype(np.int64),
                        factor=factor, fine=fine)
    print("wrote", OUT, fine.shape)


if __name__ == "__main__":
    main()
