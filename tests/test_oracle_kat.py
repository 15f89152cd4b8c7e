"""Known-answer tests of the CPU oracle, taken from the reference's own unit tests
(micromd tests/potential/*.cc, tests/misc/test_linear_hash.cc, tests/misc/test_box.cc).
CPU only."""
import ctypes as C

import numpy as np
import pytest

import bdb200 as bd


def test_softcore_known_answers(oracle):
    # tests/potential/test_softcore_potential.cc:55-76 (energy = diameter = 1, r = (0.5,0,0))
    r = [0.5, 0, 0]
    cases = {(2, 1): (0.75, 1.0), (2, 2): (0.5625, 1.5), (3, 3): (0.669921875, 1.72265625),
             (4, 4): (0.7724761962890625, 1.64794921875), (5, 5): (0.8532151877880096, 1.3761535286903381)}
    for (p, q), (e_want, fx_want) in cases.items():
        e, f = oracle.eval_potential([bd.softcore(p, q, 1.0, 1.0)], r)
        assert e == pytest.approx(e_want, rel=1e-12)
        assert f[0] == pytest.approx(fx_want, rel=1e-12)
        assert f[1] == 0 and f[2] == 0
        # zero outside the diameter (test_softcore_potential.cc:41-52)
        e, f = oracle.eval_potential([bd.softcore(p, q, 1.0, 1.0)], [10 * 0.5, 0, 0])
        assert e == 0 and not f.any()


def test_softwell_known_answers(oracle):
    # tests/potential/test_softwell_potential.cc:56-89
    r = [1.2, 3.4, 5.6]
    e, f = oracle.eval_potential([bd.softwell(2, 1.23, 4.56)], r)
    assert e == pytest.approx(-0.39255126347584784, rel=1e-12)
    np.testing.assert_allclose(f, [-0.01446003033358149, -0.040970085945147554, -0.06748014155671363], rtol=1e-12)
    e, f = oracle.eval_potential([bd.softwell(6, 1.23, 4.56)], r)
    assert e == pytest.approx(-0.11485401188866758, rel=1e-12)
    np.testing.assert_allclose(f, [-0.016901052221443922, -0.047886314627424445, -0.07887157703340497], rtol=1e-12)
    e, _ = oracle.eval_potential([bd.softwell(6, 1.23, 4.56)], [4.56, 0, 0])
    assert e == pytest.approx(-1.23 / 2)


def test_spring_family_formulas(oracle):
    # tests/potential/test_semispring_potential.cc:56-73, test_spring_potential.cc, test_harmonic_potential.cc
    k, b = 1.23, 4.56
    r = np.array([-4.0, 5.0, -6.0])
    rn = np.linalg.norm(r)
    for make in (bd.semispring, bd.spring):
        e, f = oracle.eval_potential([make(k, b)], r)
        assert e == pytest.approx(k / 2 * (rn - b) ** 2, rel=1e-12)
        np.testing.assert_allclose(f, -k * (rn - b) * r / rn, rtol=1e-12)
    # semispring is flat inside b, spring is not
    e, f = oracle.eval_potential([bd.semispring(k, b)], [1.0, 0, 0])
    assert e == 0 and not f.any()
    e, f = oracle.eval_potential([bd.spring(k, b)], [1.0, 0, 0])
    assert e == pytest.approx(k / 2 * (1 - b) ** 2) and f[0] == pytest.approx(-k * (1 - b))
    # spring force is kept finite (zero) at r = 0
    _, f = oracle.eval_potential([bd.spring(k, b)], [0.0, 0, 0])
    assert not f.any()
    e, f = oracle.eval_potential([bd.harmonic(k)], r)
    assert e == pytest.approx(k / 2 * rn**2, rel=1e-12)
    np.testing.assert_allclose(f, -k * r, rtol=1e-12)


def test_sum_scaled_cutoff_combinators(oracle):
    # tests/potential/test_sum_potential.cc, test_scaled_potential.cc, test_cutoff_potential.cc
    r = [0.1, 0.15, -0.05]
    e1, f1 = oracle.eval_potential([bd.softcore(2, 3, 2.0, 0.3)], r)
    e2, f2 = oracle.eval_potential([bd.softcore(8, 3, 1.5, 0.24)], r)
    e, f = oracle.eval_potential([bd.softcore(2, 3, 2.0, 0.3, factor=0.25), bd.softcore(8, 3, 1.5, 0.24, factor=0.75)], r)
    assert e == pytest.approx(0.25 * e1 + 0.75 * e2, rel=1e-14)
    np.testing.assert_allclose(f, 0.25 * f1 + 0.75 * f2, rtol=1e-14)
    # cutoff: zero at and beyond the cutoff distance, untouched inside
    w = bd.softwell(6, 1.0, 0.2, cutoff=0.4)
    e_in, f_in = oracle.eval_potential([w], [0.39, 0, 0])
    e_raw, f_raw = oracle.eval_potential([bd.softwell(6, 1.0, 0.2)], [0.39, 0, 0])
    assert e_in == e_raw and np.array_equal(f_in, f_raw)
    e_out, f_out = oracle.eval_potential([w], [0.4, 0, 0])
    assert e_out == 0 and not f_out.any()


def test_linear_hash_known_answers(oracle):
    # tests/misc/test_linear_hash.cc:23-59
    lib = oracle.lib
    lib.orc_linear_hash.restype = C.c_uint32
    lib.orc_linear_hash.argtypes = [C.POINTER(C.c_uint32), C.c_uint32, C.c_uint32, C.c_uint32, C.c_uint32]

    def h(coeff, mod, x, y, z):
        return lib.orc_linear_hash((C.c_uint32 * 3)(*coeff), mod, x, y, z)

    c = (2, 3, 5)
    assert [h(c, 11, *v) for v in [(1, 0, 0), (0, 1, 0), (0, 0, 1), (1, 1, 1)]] == [2, 3, 5, 10]
    want = {(1000, 10, 100): 0, (1000, 100, 10): 7, (10, 1000, 100): 0, (100, 1000, 10): 5,
            (10, 100, 1000): 7, (100, 10, 1000): 5}
    for v, w in want.items():
        assert h(c, 11, *v) == w
    big = (3416542937, 3481543853, 3912176573)  # no 32-bit wraparound
    assert h(big, 17, 1234, 5678, 9012) == 14
    assert h(big, 17, 3456, 7890, 1234) == 4
    assert h(big, 17, 5678, 9012, 3456) == 0
    d = (3929498747, 1008281837, 1832832077)
    assert h(d, 1021, 1021, 1021, 1021) == 0


def _brute_pairs(x, dcut, period=None):
    d = x[:, None, :] - x[None, :, :]
    if period is not None:
        L = np.asarray(period)
        d = d - np.rint(d * (1 / L)) * L
    r2 = d[..., 0] * d[..., 0] + d[..., 1] * d[..., 1] + d[..., 2] * d[..., 2]
    i, j = np.nonzero(np.triu(r2 <= dcut * dcut, k=1))
    return set(zip(i.tolist(), j.tolist()))


@pytest.mark.parametrize("kind,period", [(bd.BOX_OPEN, None), (bd.BOX_PERIODIC, (0.9, 1.0, 1.1))])
def test_neighbor_search_equals_brute_force(oracle, kind, period):
    # tests/misc/test_neighbor_searcher.cc:54-127: 100 N(0,1) points, dcut 0.3; the result as a
    # multiset must equal the brute-force set (no duplicates, nothing missing)
    x = np.random.default_rng(0).normal(size=(100, 3))
    got = oracle.nsearch_pairs(kind, period, 100, 0.3, x)
    as_list = [tuple(p) for p in got.tolist()]
    assert len(as_list) == len(set(as_list))
    assert set(as_list) == _brute_pairs(x, 0.3, period)
    assert all(i < j for i, j in as_list)


def test_neighbor_search_edge_cases(oracle):
    # empty input, one point, coincident points, tiny periodic box (one bin per axis)
    assert len(oracle.nsearch_pairs(bd.BOX_OPEN, None, 0, 0.3, np.zeros((0, 3)))) == 0
    assert len(oracle.nsearch_pairs(bd.BOX_OPEN, None, 1, 0.3, np.zeros((1, 3)))) == 0
    x = np.zeros((5, 3))
    assert len(oracle.nsearch_pairs(bd.BOX_OPEN, None, 5, 0.3, x)) == 10
    x = np.random.default_rng(3).uniform(-3, 3, size=(40, 3))
    got = oracle.nsearch_pairs(bd.BOX_PERIODIC, (0.5, 0.5, 0.5), 40, 0.3, x)
    as_list = [tuple(p) for p in got.tolist()]
    assert len(as_list) == len(set(as_list))
    assert set(as_list) == _brute_pairs(x, 0.3, (0.5, 0.5, 0.5))


def test_normal_stream_statistics(oracle):
    z = oracle.normals(2024, 400_000)
    assert abs(z.mean()) < 5e-3 and abs(z.std() - 1) < 5e-3
    assert abs((z**4).mean() - 3) < 0.05
    assert np.array_equal(z[:1000], oracle.normals(2024, 1000))  # reproducible prefix
