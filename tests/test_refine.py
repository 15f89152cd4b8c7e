"""Spline refinement between the 10Mb and the 100kb phase.

tests/golden/refine_spline.npz was produced by the reference's own refinement module
(5-sim-genome/src/refine/refinement.py, scipy splprep/splev) via tests/golden/make_refine_golden.py.
The oracle's C restatement (dense solve) is pinned to it on the CPU; the engine (closed-form not-a-knot
elimination on the device) is compared with both on the GPU.  Tolerance: 1e-10 relative to the extent of the
conformation -- FITPACK, the dense solve and the tridiagonal solve agree to rounding.
"""
import os

import numpy as np
import pytest

import bdb200 as bd

HERE = os.path.dirname(os.path.abspath(__file__))
TOL = 1e-10


def golden():
    g = np.load(os.path.join(HERE, "golden", "refine_spline.npz"))
    return {k: g[k] for k in g.files}


def run(backend, g, **kw):
    return backend.refine_spline(g["coarse"], g["coarse_ranges"], int(g["factor"]), g["fine_ranges"], len(g["fine"]),
                                 attach=g["attach"], **kw)


def err(a, b):
    return np.abs(a - b).max() / np.abs(b).max()


def test_golden_is_the_reference():
    """Regenerate with the reference's module wherever it (and scipy) exists."""
    if not os.path.exists("/root/reference/5-sim-genome/src/refine/refinement.py"):
        pytest.skip("no /root/reference here")
    pytest.importorskip("scipy")
    import importlib.util

    spec = importlib.util.spec_from_file_location("reference_refinement", "/root/reference/5-sim-genome/src/refine/refinement.py")
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    g = golden()
    for (b, e), (fb, fe) in zip(g["coarse_ranges"], g["fine_ranges"]):
        chain = mod.refine_path_spline(g["coarse"][b:e], (e - b) * int(g["factor"]))
        assert np.array_equal(chain[: fe - fb], g["fine"][fb:fe])


def test_oracle_matches_the_reference_spline(oracle):
    g = golden()
    out = run(oracle, g)
    assert err(out, g["fine"]) <= TOL
    # the interpolant passes through the coarse points: bin midpoints f = factor * i + (factor - 1) / 2 exist
    # for odd factors only, so check with factor 1 (fine == coarse)
    same = oracle.refine_spline(g["coarse"], g["coarse_ranges"], 1, g["coarse_ranges"], len(g["coarse"]))
    assert err(same, g["coarse"]) <= 1e-13
    with pytest.raises(bd.BackendError):
        oracle.refine_spline(g["coarse"][:3], [[0, 3]], 10, [[0, 30]], 30)


@pytest.mark.gpu
def test_engine_matches_reference_and_oracle(engine, oracle):
    g = golden()
    out = run(engine, g)
    assert err(out, g["fine"]) <= TOL and err(out, run(oracle, g)) <= TOL
    # rows no chain covers keep the caller's values; attached beads copy their NOR bead
    init = np.full((len(g["fine"]) + 5, 3), 7.0)
    out2 = engine.refine_spline(g["coarse"], g["coarse_ranges"], int(g["factor"]), g["fine_ranges"], len(init),
                                attach=g["attach"], fine_init=init)
    assert np.array_equal(out2[-5:], init[-5:]) and np.array_equal(out2[: len(out)], out)
    for nor, nuc in g["attach"]:
        assert np.array_equal(out[nuc], out[nor])
    same = engine.refine_spline(g["coarse"], g["coarse_ranges"], 1, g["coarse_ranges"], len(g["coarse"]))
    assert err(same, g["coarse"]) <= 1e-13
    # a long chain (the tridiagonal sweep) against the oracle's dense solve
    rng = np.random.default_rng(3)
    long_chain = np.cumsum(rng.normal(size=(300, 3)), axis=0)
    a = engine.refine_spline(long_chain, [[0, 300]], 7, [[0, 2100]], 2100)
    b = oracle.refine_spline(long_chain, [[0, 300]], 7, [[0, 2100]], 2100)
    assert err(a, b) <= TOL
    for bad in ([[0, 3]], [[0, 400]]):
        with pytest.raises(bd.BackendError):
            engine.refine_spline(long_chain, bad, 7, [[0, 10]], 2100)
    with pytest.raises(bd.BackendError):
        engine.refine_spline(long_chain, [[0, 300]], 7, [[0, 2101]], 2101)   # longer than factor x coarse
