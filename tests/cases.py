"""Small systems shared by the parity tests (CPU oracle pin, golden fixtures, GPU parity)."""
import numpy as np

import bdb200 as bd
import scenarios as sc

SMALL = {
    "ab_box": lambda: sc.ab_box(20, 20),
    "ab_sphere": lambda: sc.ab_sphere(20, 20, inner_wall=True),
    "genome_10mb": lambda: sc.genome_10mb(bend_energy=0.7),
    "genome_100kb": lambda: sc.genome_100kb(3000, 30, bead_scale=0.8, bond_scale=0.9),
    "chromatin_1kb": lambda: sc.chromatin_1kb(3, 300),
}

GOLDEN = {
    "ab_box": lambda: sc.ab_box(10, 20),
    "ab_sphere": lambda: sc.ab_sphere(10, 20, inner_wall=True),
    "genome_10mb": lambda: sc.genome_10mb(total=320, bend_energy=0.7),
    "genome_100kb": lambda: sc.genome_100kb(600, 12, bead_scale=0.8, bond_scale=0.9),
    "chromatin_1kb": lambda: sc.chromatin_1kb(2, 150),
}
GOLDEN_STEPS = 25
GOLDEN_SEED = 11


def build(backend, scenario, n_replicas=1, replicas=None):
    s = bd.System(backend, scenario.n, n_replicas)
    ids = scenario.apply(s, replicas)
    return s, ids


def rel_force_error(f, f_ref):
    """max_i |f_i - f_ref_i| / max_i |f_ref_i|  (the 'relative to the largest force' convention
    used for every force tolerance in this repository)"""
    scale = np.linalg.norm(f_ref, axis=1).max()
    return np.linalg.norm(f - f_ref, axis=1).max() / (scale if scale > 0 else 1.0)


def brute_pairs(x, dcut, period=None):
    """All unordered pairs within dcut with the reference's fp64 arithmetic
    (md::periodic_box::shortest_displacement, squared_norm left to right)."""
    x = np.asarray(x)
    out = []
    L = None if period is None else np.asarray(period, dtype=np.float64)
    for i in range(len(x) - 1):
        d = x[i] - x[i + 1:]
        if L is not None:
            d = d - np.rint(d * (1 / L)) * L
        r2 = d[:, 0] * d[:, 0] + d[:, 1] * d[:, 1] + d[:, 2] * d[:, 2]
        j = np.nonzero(r2 <= dcut * dcut)[0] + i + 1
        out.append(np.stack([np.full(len(j), i), j], axis=1))
    return np.concatenate(out).astype(np.uint32) if out else np.zeros((0, 2), np.uint32)


def sort_pairs(p):
    p = np.asarray(p, dtype=np.uint32).reshape(-1, 2)
    p = np.stack([p.min(axis=1), p.max(axis=1)], axis=1)
    order = np.lexsort((p[:, 1], p[:, 0]))
    return p[order]


# ---- the sizes BASELINE.json names (N = 4 000, 640 x 1 024 replicas, 62 600, 10^6) ----------

NAMED = {
    "ab_box_4000": lambda: sc.ab_box(200, 20),                       # C1
    "ab_sphere_4000": lambda: sc.ab_sphere(200, 20),                 # C2
    "genome_10mb_640": lambda: sc.genome_10mb(640, bend_energy=0.7),  # C3 (one of the 1 024 replicas)
    "genome_100kb_62600": lambda: sc.genome_100kb(62000, 300),       # C4
    "chromatin_1kb_1M": lambda: sc.chromatin_1kb(100, 10000),        # C5 at 10^6 beads
}
NAMED_STEPS = 10
NAMED_SEED = 23
NAMED_SAMPLE = 1000


def sample_beads(n):
    """the beads whose forces / positions the full-size fixtures store (fixed, spread over the chains)"""
    return np.unique(np.linspace(0, n - 1, min(n, NAMED_SAMPLE)).astype(np.int64))


def digest(a):
    """sha256 of an array's bytes: bit-exact identity of a full-size result"""
    import hashlib

    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def pair_digest(pairs):
    """order-independent identity of a pair set: count, sum and xor of the 64-bit keys i << 32 | j"""
    p = sort_pairs(pairs).astype(np.uint64)
    key = (p[:, 0] << np.uint64(32)) | p[:, 1]
    return np.array([len(key), int(key.sum(dtype=np.uint64)), int(np.bitwise_xor.reduce(key)) if len(key) else 0],
                    dtype=np.uint64)
