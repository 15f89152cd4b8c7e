"""ctypes binding of the Brownian-dynamics C ABI (include/bd_cuda.h, include/bd_api_decl.inc).

The same entry points exist with three prefixes:

* ``bd_``  -- the B200 engine, ``2022a-genome-dynamics_b200/libbd_b200.so`` (the product);
* ``orc_`` -- the CPU restatement, ``oracle/liboracle.so`` (test infrastructure only);
* ``ref_`` -- the unmodified reference micromd, ``oracle/_ref/libmicromd_ref.so`` (test only).

``load_engine()`` is the product path.  It raises if the CUDA library is missing: there is no CPU
fallback, and this module does not know where the checkers live -- their loaders are test
infrastructure in ``oracle/checkers.py`` and construct ``Backend(path, prefix)`` themselves.
"""
from __future__ import annotations

import ctypes as C
import os
from typing import Optional, Sequence

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)

# ---- include/bd_types.h -----------------------------------------------------------------------

BOX_OPEN, BOX_PERIODIC = 0, 1
TERM_NONE, TERM_SOFTCORE, TERM_SOFTWELL, TERM_HARMONIC, TERM_SPRING, TERM_SEMISPRING = range(6)
FF_ALL, FF_PAIR, FF_BOND, FF_BEND, FF_WALL, FF_POINT = range(6)
WALL_SPHERE, WALL_ELLIPSOID = 0, 1
WALL_INWARD, WALL_OUTWARD = 0, 1
MAX_TERMS, MAX_TYPES = 4, 16

TERM_DTYPE = np.dtype(
    [
        ("kind", "<i4"),
        ("p", "<i4"),
        ("q", "<i4"),
        ("reserved", "<i4"),
        ("factor", "<f8"),
        ("energy", "<f8"),
        ("length", "<f8"),
        ("cutoff", "<f8"),
    ]
)
assert TERM_DTYPE.itemsize == 48


class StepConfig(C.Structure):
    _fields_ = [
        ("temperature", C.c_double),
        ("timestep", C.c_double),
        ("spacestep", C.c_double),
        ("steps", C.c_int64),
        ("seed", C.c_uint64),
        ("first_step", C.c_int64),
        ("replay_normals", C.c_void_p),
    ]


class Counters(C.Structure):
    _fields_ = [
        ("steps_done", C.c_int64),
        ("list_rebuilds", C.c_int64),
        ("pair_interactions", C.c_int64),
        ("listed_pairs", C.c_int64),
        ("overflow_beads", C.c_int64),
        ("reserved", C.c_int64 * 3),
    ]


def term(kind, p=0, q=0, factor=1.0, energy=0.0, length=0.0, cutoff=0.0):
    return (kind, p, q, 0, factor, energy, length, cutoff)


def softcore(p, q, energy, diameter, factor=1.0, cutoff=0.0):
    """factor * md::softcore_potential<p,q>{energy, diameter}"""
    return term(TERM_SOFTCORE, p, q, factor, energy, diameter, cutoff)


def softwell(p, energy, decay, factor=1.0, cutoff=0.0):
    """factor * [apply_cutoff](md::softwell_potential<p>{energy, decay})"""
    return term(TERM_SOFTWELL, p, 0, factor, energy, decay, cutoff)


def harmonic(k, factor=1.0):
    return term(TERM_HARMONIC, 0, 0, factor, k, 0.0)


def spring(k, b, factor=1.0):
    return term(TERM_SPRING, 0, 0, factor, k, b)


def semispring(k, b, factor=1.0):
    return term(TERM_SEMISPRING, 0, 0, factor, k, b)


def none_term():
    return term(TERM_NONE)


def table(rows: Sequence[Sequence[tuple]]) -> np.ndarray:
    """rows[cell] = list of terms; all cells padded to the same number of terms with NONE."""
    nterms = max(len(r) for r in rows)
    out = np.zeros((len(rows), nterms), dtype=TERM_DTYPE)
    for i, r in enumerate(rows):
        for k, t in enumerate(r):
            out[i, k] = t
    return out


def pair_table(ntypes: int, fn) -> np.ndarray:
    """fn(ti, tj) -> list of terms; returns the [ntypes*ntypes, nterms] table."""
    return table([fn(ti, tj) for ti in range(ntypes) for tj in range(ntypes)])


def field_table(ntypes: int, fn) -> np.ndarray:
    return table([fn(ti) for ti in range(ntypes)])


# ---- library loading ---------------------------------------------------------------------------

_P = C.POINTER
_dbl_p = _P(C.c_double)
_u8_p = _P(C.c_uint8)
_u32_p = _P(C.c_uint32)


def _api(prefix: str):
    i32, i64, vp, dbl = C.c_int32, C.c_int64, C.c_void_p, C.c_double
    return {
        "create": [i64, i32, i32, _P(vp)],
        "destroy": [vp],
        "set_box": [vp, i32, _dbl_p],
        "set_types": [vp, _u8_p, i32, _dbl_p],
        "set_positions": [vp, i32, _dbl_p],
        "get_positions": [vp, i32, _dbl_p],
        "add_pair_ff": [vp, i32, vp, dbl, _u8_p, _P(i32)],
        "update_pair_ff": [vp, i32, vp, dbl],
        "add_bond_ff": [vp, i32, vp, _u32_p, i64, _P(i32)],
        "update_bond_ff": [vp, i32, vp],
        "set_bond_pairs": [vp, i32, _u32_p, i64],
        "set_bond_min_image": [vp, i32, i32],
        "add_bend_ff": [vp, _u32_p, _dbl_p, i64, _P(i32)],
        "add_wall_ff": [vp, i32, i32, i32, vp, _dbl_p, _dbl_p, _P(i32)],
        "update_wall_ff": [vp, i32, vp, _dbl_p, _dbl_p],
        "get_wall_stats": [vp, i32, i32, _dbl_p],
        "add_point_ff": [vp, i32, vp, _dbl_p, _u8_p, _P(i32)],
        "set_ff_enabled": [vp, i32, i32, i32],
        "set_external_forces": [vp, i32, _dbl_p],
        "compute_forces": [vp, i32, i32, i32, _dbl_p],
        "compute_energy": [vp, i32, i32, i32, _dbl_p],
        "step": [vp, _P(StepConfig)],
        "get_last_timestep": [vp, i32, _dbl_p],
        "get_neighbor_list": [vp, i32, _u32_p, i64, _P(i64)],
        "get_counters": [vp, _P(Counters)],
        "nsearch_buckets": [i32, _dbl_p, i64, dbl, _dbl_p, i64, _u32_p, i32],
        "nsearch_pairs": [i32, _dbl_p, i64, dbl, _dbl_p, i64, _u32_p, i64, _P(i64), i32],
    }


ENGINE_ONLY = {
    "set_stream": [C.c_void_p, C.c_void_p],
    "sync": [C.c_void_p],
    "set_option": [C.c_void_p, C.c_char_p, C.c_double],
    "get_option": [C.c_void_p, C.c_char_p, _dbl_p],
    "set_positions_device": [C.c_void_p, C.c_void_p],
    "get_positions_device": [C.c_void_p, C.c_void_p],
    "get_launch_count": [C.c_void_p, _P(C.c_int64)],
    "update_wall_ff_replica": [C.c_void_p, C.c_int32, C.c_int32, _dbl_p, _dbl_p],
    "get_wall_geometry": [C.c_void_p, C.c_int32, C.c_int32, _dbl_p, _dbl_p],
    "get_wall_stats_all": [C.c_void_p, C.c_int32, _dbl_p],
    "set_wall_relaxation": [C.c_void_p, _P(C.c_int32), C.c_int32, _dbl_p, C.c_double],
    "noise_statistics": [C.c_uint64, C.c_int64, C.c_int64, C.c_int64, C.c_int32, C.c_int32, _dbl_p, C.c_int32],
    "set_all_positions": [C.c_void_p, C.c_void_p],
    "nsearch_query": [C.c_int32, _dbl_p, C.c_double, _dbl_p, C.c_int64, _dbl_p, C.c_int64, _P(C.c_int64), _u32_p, C.c_int64,
                      C.c_int32],
    "nsearch_pairs_ordered": [C.c_int32, _dbl_p, C.c_int64, C.c_double, _dbl_p, C.c_int64, _u32_p, C.c_int64, _P(C.c_int64),
                              C.c_int32],
    "glue_update": [C.c_void_p, C.c_int32, C.c_int32, C.c_double, C.c_double, C.c_double, C.c_int64, C.c_uint64, C.c_int64,
                    C.c_int32, _P(C.c_int64), _P(C.c_int64), _P(C.c_int64)],
    "get_bond_pairs": [C.c_void_p, C.c_int32, _u32_p, C.c_int64, _P(C.c_int64)],
    "get_all_positions": [C.c_void_p, C.c_void_p],
    "get_all_positions_async": [C.c_void_p, C.c_void_p],
    "set_owned": [C.c_void_p, _u8_p],
    "set_owned_device": [C.c_void_p, C.c_void_p],
    "gather_positions_indexed": [C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p],
    "displacement_pointer": [C.c_void_p, _P(C.c_void_p)],
    "scatter_positions_indexed": [C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p],
    "contact_update": [C.c_void_p, C.c_int32, C.c_double],
    "contact_accumulate": [C.c_void_p, C.c_int32, C.POINTER(C.c_uint32), C.c_int64, C.POINTER(C.c_int64)],
    "contact_clear": [C.c_void_p, C.c_int32],
    "refine_spline": [_dbl_p, _P(C.c_int64), C.c_int64, C.c_int32, _P(C.c_int64), _P(C.c_int64), C.c_int64, C.c_int64,
                      C.c_int64, _dbl_p, C.c_int32],
    "search_pairs": [C.c_void_p, C.c_int32, C.c_double, C.c_int32, _u32_p, C.c_int64, C.POINTER(C.c_int64)],
    "pair_distances": [C.c_void_p, C.c_int32, _u32_p, C.c_int64, C.c_int32, _dbl_p],
    "snapshot_begin": [C.c_void_p, C.c_int32, C.c_int32],
    "snapshot_wait": [C.c_void_p, C.POINTER(C.POINTER(C.c_float)), C.POINTER(C.c_int64)],
}


class BackendError(RuntimeError):
    pass


class Backend:
    """One shared library exposing the API under `prefix`."""

    def __init__(self, path: str, prefix: str):
        if not os.path.exists(path):
            raise FileNotFoundError(path)
        self.path = path
        self.prefix = prefix
        self.lib = C.CDLL(path, mode=C.RTLD_GLOBAL if prefix == "bd_" else C.RTLD_LOCAL)
        self._last_error = getattr(self.lib, prefix + "last_error")
        self._last_error.restype = C.c_char_p
        self._last_error.argtypes = []
        self.fn = {}
        for name, argtypes in _api(prefix).items():
            f = getattr(self.lib, prefix + name)
            f.restype = C.c_int
            f.argtypes = argtypes
            self.fn[name] = f
        if prefix == "bd_":
            for name, argtypes in ENGINE_ONLY.items():
                f = getattr(self.lib, prefix + name)
                f.restype = C.c_int
                f.argtypes = argtypes
                self.fn[name] = f
        else:
            for name, argtypes in {
                "normals": [C.c_uint64, C.c_int64, _dbl_p],
                "sfc64": [C.c_uint64, C.c_int64, _P(C.c_uint64)],
                "eval_potential": [C.c_void_p, C.c_int32, _dbl_p, _dbl_p, _dbl_p],
                "quantize_positions": [_dbl_p, C.c_int64, C.c_int32, _P(C.c_float)],  # oracle only
                "refine_spline": [_dbl_p, _P(C.c_int64), C.c_int64, C.c_int32, _P(C.c_int64), _P(C.c_int64), C.c_int64,
                                  C.c_int64, C.c_int64, _dbl_p],  # oracle only
            }.items():
                if name in ("quantize_positions", "refine_spline") and prefix != "orc_":
                    continue
                f = getattr(self.lib, prefix + name)
                f.restype = C.c_int
                f.argtypes = argtypes
                self.fn[name] = f

    def call(self, name: str, *args):
        rc = self.fn[name](*args)
        if rc != 0:
            msg = self._last_error()
            raise BackendError(f"{self.prefix}{name} failed ({rc}): {msg.decode() if msg else ''}")

    # -- stateless helpers --------------------------------------------------------------------

    def nsearch_buckets(self, box_kind, period, particle_count, spacing, xyz, device=0):
        xyz = np.ascontiguousarray(xyz, dtype=np.float64)
        period = np.ascontiguousarray(period if period is not None else [1, 1, 1], dtype=np.float64)
        out = np.zeros(len(xyz), dtype=np.uint32)
        self.call("nsearch_buckets", box_kind, _dp(period), int(particle_count), float(spacing),
                  _dp(xyz), len(xyz), out.ctypes.data_as(_u32_p), device)
        return out

    def nsearch_pairs(self, box_kind, period, particle_count, dcut, xyz, cap=None, device=0, ordered=False):
        """ordered (engine only): the reference's emission order (orc_ / ref_ always emit in it)"""
        fn = "nsearch_pairs_ordered" if ordered else "nsearch_pairs"
        xyz = np.ascontiguousarray(xyz, dtype=np.float64)
        period = np.ascontiguousarray(period if period is not None else [1, 1, 1], dtype=np.float64)
        cap = int(cap if cap is not None else max(1024, 64 * len(xyz)))
        while True:
            out = np.zeros((cap, 2), dtype=np.uint32)
            count = C.c_int64(0)
            self.call(fn, box_kind, _dp(period), int(particle_count), float(dcut),
                      _dp(xyz), len(xyz), out.ctypes.data_as(_u32_p), cap, C.byref(count), device)
            if count.value <= cap:
                return out[: count.value]
            cap = count.value

    def quantize_positions(self, xyz, fraction_bits: int = 16) -> np.ndarray:
        xyz = np.ascontiguousarray(xyz, dtype=np.float64).reshape(-1, 3)
        out = np.empty(xyz.shape, dtype=np.float32)
        self.call("quantize_positions", _dp(xyz), len(xyz), int(fraction_bits), out.ctypes.data_as(_P(C.c_float)))
        return out

    def nsearch_query(self, box_kind, period, dcut, xyz, queries, device=0):
        """md::neighbor_searcher::query for a batch of points: (start[m + 1], index[total]) (bd_cuda.h)"""
        xyz = np.ascontiguousarray(xyz, dtype=np.float64).reshape(-1, 3)
        q = np.ascontiguousarray(queries, dtype=np.float64).reshape(-1, 3)
        period = np.ascontiguousarray(period if period is not None else [1, 1, 1], dtype=np.float64)
        start = np.zeros(len(q) + 1, dtype=np.int64)
        self.call("nsearch_query", box_kind, _dp(period), float(dcut), _dp(xyz), len(xyz), _dp(q), len(q),
                  start.ctypes.data_as(_P(C.c_int64)), None, 0, device)
        index = np.zeros(max(int(start[-1]), 1), dtype=np.uint32)
        self.call("nsearch_query", box_kind, _dp(period), float(dcut), _dp(xyz), len(xyz), _dp(q), len(q),
                  start.ctypes.data_as(_P(C.c_int64)), index.ctypes.data_as(_u32_p), int(start[-1]), device)
        return start, index[: int(start[-1])]

    def noise_statistics(self, seed, first_step, steps, beads, replica=0, device=0) -> dict:
        """moments, tail counts and correlations of the step kernel's production noise (bd_cuda.h)"""
        out = np.zeros(20)
        self.call("noise_statistics", int(seed), int(first_step), int(steps), int(beads), int(replica), int(device),
                  _dp(out), 20)
        return {"count": out[0], "sum": out[1:6], "max_abs": out[6], "tail": out[7:14], "lag1": out[14],
                "cross": out[15:18], "sq_cross": out[18], "triples": out[19]}

    def refine_spline(self, coarse_xyz, coarse_ranges, factor, fine_ranges, n_fine, attach=None, fine_init=None,
                      device=0) -> np.ndarray:
        """5-sim-genome/src/refine (spline refinement of every chain + nucleolar attachment); [n_fine, 3]."""
        coarse = np.ascontiguousarray(coarse_xyz, dtype=np.float64).reshape(-1, 3)
        cr = np.ascontiguousarray(coarse_ranges, dtype=np.int64).reshape(-1, 2)
        fr = np.ascontiguousarray(fine_ranges, dtype=np.int64).reshape(-1, 2)
        at = np.ascontiguousarray(attach if attach is not None else np.zeros((0, 2)), dtype=np.int64).reshape(-1, 2)
        out = np.zeros((int(n_fine), 3)) if fine_init is None else np.array(fine_init, dtype=np.float64).reshape(int(n_fine), 3)
        i64p = _P(C.c_int64)
        args = [_dp(coarse), cr.ctypes.data_as(i64p), len(cr), int(factor), fr.ctypes.data_as(i64p),
                at.ctypes.data_as(i64p), len(at), len(coarse), int(n_fine), _dp(out)]
        if self.prefix == "bd_":
            args.append(int(device))
        self.call("refine_spline", *args)
        return out

    def normals(self, seed: int, count: int) -> np.ndarray:
        out = np.zeros(count, dtype=np.float64)
        self.call("normals", seed, count, _dp(out))
        return out

    def sfc64(self, seed: int, count: int) -> np.ndarray:
        out = np.zeros(count, dtype=np.uint64)
        self.call("sfc64", seed, count, out.ctypes.data_as(_P(C.c_uint64)))
        return out

    def eval_potential(self, terms, r):
        terms = np.ascontiguousarray(np.array(terms, dtype=TERM_DTYPE).reshape(-1))
        r = np.ascontiguousarray(r, dtype=np.float64)
        e = C.c_double(0)
        f = np.zeros(3)
        self.call("eval_potential", terms.ctypes.data, len(terms), _dp(r), C.byref(e), _dp(f))
        return e.value, f


def _dp(a: np.ndarray):
    return a.ctypes.data_as(_dbl_p)


def _mask(mask, n) -> Optional[np.ndarray]:
    if mask is None:
        return None
    m = np.ascontiguousarray(mask, dtype=np.uint8)
    assert m.shape == (n,)
    return m


class System:
    """md::system + its registered forcefields, living in one backend."""

    def __init__(self, backend: Backend, n_beads: int, n_replicas: int = 1, device: int = 0):
        self.b = backend
        self.n = int(n_beads)
        self.replicas = int(n_replicas)
        self.ntypes = 1
        self.h = C.c_void_p()
        backend.call("create", self.n, self.replicas, device, C.byref(self.h))

    def close(self):
        if self.h:
            self.b.call("destroy", self.h)
            self.h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # -- state ------------------------------------------------------------------------------
    def set_box(self, kind, period=(1.0, 1.0, 1.0)):
        p = np.ascontiguousarray(period, dtype=np.float64)
        self.b.call("set_box", self.h, kind, _dp(p))

    def set_types(self, types, mobility):
        t = np.ascontiguousarray(types, dtype=np.uint8)
        m = np.ascontiguousarray(mobility, dtype=np.float64)
        assert t.shape == (self.n,)
        self.ntypes = len(m)
        self.b.call("set_types", self.h, t.ctypes.data_as(_u8_p), len(m), _dp(m))

    def set_positions(self, xyz, replica=0):
        x = np.ascontiguousarray(xyz, dtype=np.float64)
        assert x.shape == (self.n, 3)
        self.b.call("set_positions", self.h, replica, _dp(x))

    def get_positions(self, replica=0):
        x = np.zeros((self.n, 3), dtype=np.float64)
        self.b.call("get_positions", self.h, replica, _dp(x))
        return x

    # -- forcefields ------------------------------------------------------------------------
    def _table(self, tab, cells):
        tab = np.ascontiguousarray(tab, dtype=TERM_DTYPE)
        assert tab.ndim == 2 and tab.shape[0] == cells, (tab.shape, cells)
        return tab

    def add_pair_ff(self, tab, dcut, targets=None):
        tab = self._table(tab, self.ntypes**2)
        m = _mask(targets, self.n)
        ff = C.c_int32(-1)
        self.b.call("add_pair_ff", self.h, tab.shape[1], tab.ctypes.data, float(dcut),
                    m.ctypes.data_as(_u8_p) if m is not None else None, C.byref(ff))
        return ff.value

    def update_pair_ff(self, ff, tab, dcut):
        tab = self._table(tab, self.ntypes**2)
        self.b.call("update_pair_ff", self.h, ff, tab.ctypes.data, float(dcut))

    def add_bond_ff(self, tab, pairs):
        tab = self._table(tab, self.ntypes**2)
        p = np.ascontiguousarray(pairs, dtype=np.uint32).reshape(-1, 2)
        ff = C.c_int32(-1)
        self.b.call("add_bond_ff", self.h, tab.shape[1], tab.ctypes.data,
                    p.ctypes.data_as(_u32_p), len(p), C.byref(ff))
        return ff.value

    def set_bond_pairs(self, ff, pairs):
        """Replace the pair list of a bond forcefield (a dynamic glue / loop list; may be empty)."""
        p = np.ascontiguousarray(pairs, dtype=np.uint32).reshape(-1, 2)
        self.b.call("set_bond_pairs", self.h, ff, p.ctypes.data_as(_u32_p), len(p))

    def set_bond_min_image(self, ff, on=True):
        self.b.call("set_bond_min_image", self.h, ff, int(bool(on)))

    def update_bond_ff(self, ff, tab):
        tab = self._table(tab, self.ntypes**2)
        self.b.call("update_bond_ff", self.h, ff, tab.ctypes.data)

    def add_bend_ff(self, triples, energies):
        t = np.ascontiguousarray(triples, dtype=np.uint32).reshape(-1, 3)
        e = np.ascontiguousarray(energies, dtype=np.float64)
        assert len(e) == len(t)
        ff = C.c_int32(-1)
        self.b.call("add_bend_ff", self.h, t.ctypes.data_as(_u32_p), _dp(e), len(t), C.byref(ff))
        return ff.value

    def add_wall_ff(self, shape, direction, tab, center, semiaxes):
        tab = self._table(tab, self.ntypes)
        c = np.ascontiguousarray(center, dtype=np.float64)
        s = np.ascontiguousarray(semiaxes, dtype=np.float64)
        ff = C.c_int32(-1)
        self.b.call("add_wall_ff", self.h, shape, direction, tab.shape[1], tab.ctypes.data,
                    _dp(c), _dp(s), C.byref(ff))
        return ff.value

    def update_wall_ff(self, ff, center, semiaxes, tab=None, replica=None):
        """geometry (and optionally the table) of a wall: of every replica, or of one (engine only).
        center = semiaxes = None changes the table only (engine only)."""
        if tab is not None:
            tab = self._table(tab, self.ntypes)
        if center is None and semiaxes is None:
            self.b.call("update_wall_ff", self.h, ff, tab.ctypes.data, None, None)
            return
        c = np.ascontiguousarray(center, dtype=np.float64)
        s = np.ascontiguousarray(semiaxes, dtype=np.float64)
        if replica is not None:
            assert tab is None
            self.b.call("update_wall_ff_replica", self.h, ff, replica, _dp(c), _dp(s))
            return
        self.b.call("update_wall_ff", self.h, ff, tab.ctypes.data if tab is not None else None,
                    _dp(c), _dp(s))

    def get_wall_geometry(self, ff, replica=0):
        c, s = np.zeros(3), np.zeros(3)
        self.b.call("get_wall_geometry", self.h, ff, replica, _dp(c), _dp(s))
        return c, s

    def get_wall_stats_all(self, ff):
        out = np.zeros((self.replicas, 4))
        self.b.call("get_wall_stats_all", self.h, ff, _dp(out))
        return out

    def set_wall_relaxation(self, ffs, spring, mobility):
        """device-resident semiaxis feedback of the walls `ffs` (empty = off), bd_cuda.h"""
        f = (C.c_int32 * max(len(ffs), 1))(*ffs)
        sp = np.ascontiguousarray(spring, dtype=np.float64)
        self.b.call("set_wall_relaxation", self.h, f, len(ffs), _dp(sp), float(mobility))

    def get_wall_stats(self, ff, replica=0):
        out = np.zeros(4)
        self.b.call("get_wall_stats", self.h, ff, replica, _dp(out))
        return out

    def add_point_ff(self, tab, source, targets=None):
        tab = self._table(tab, self.ntypes)
        s = np.ascontiguousarray(source, dtype=np.float64)
        m = _mask(targets, self.n)
        ff = C.c_int32(-1)
        self.b.call("add_point_ff", self.h, tab.shape[1], tab.ctypes.data, _dp(s),
                    m.ctypes.data_as(_u8_p) if m is not None else None, C.byref(ff))
        return ff.value

    def set_external_forces(self, xyz, replica=0):
        """force array of host-evaluated user forcefields, added to every whole-system evaluation; None clears"""
        if xyz is None:
            self.b.call("set_external_forces", self.h, replica, None)
            return
        f = np.ascontiguousarray(xyz, dtype=np.float64)
        assert f.shape == (self.n, 3)
        self.b.call("set_external_forces", self.h, replica, _dp(f))

    def set_ff_enabled(self, ff_class, ff, enabled):
        self.b.call("set_ff_enabled", self.h, ff_class, ff, int(bool(enabled)))

    # -- evaluation -------------------------------------------------------------------------
    def compute_forces(self, replica=0, ff_class=FF_ALL, ff=-1):
        out = np.zeros((self.n, 3), dtype=np.float64)
        self.b.call("compute_forces", self.h, replica, ff_class, ff, _dp(out))
        return out

    def compute_energy(self, replica=0, ff_class=FF_ALL, ff=-1):
        e = C.c_double(0)
        self.b.call("compute_energy", self.h, replica, ff_class, ff, C.byref(e))
        return e.value

    def step(self, steps, timestep, temperature=1.0, seed=0, first_step=1, spacestep=0.0,
             replay_normals=None):
        cfg = StepConfig()
        cfg.temperature = temperature
        cfg.timestep = timestep
        cfg.spacestep = spacestep
        cfg.steps = steps
        cfg.seed = seed
        cfg.first_step = first_step
        keep = None
        if replay_normals is not None:
            keep = np.ascontiguousarray(replay_normals, dtype=np.float64)
            assert keep.size == steps * self.replicas * self.n * 3, keep.shape
            cfg.replay_normals = keep.ctypes.data
        self.b.call("step", self.h, C.byref(cfg))
        del keep

    def get_last_timestep(self, replica=0):
        v = C.c_double(0)
        self.b.call("get_last_timestep", self.h, replica, C.byref(v))
        return v.value

    def get_neighbor_list(self, replica=0):
        cap = max(1024, 64 * self.n)
        while True:
            out = np.zeros((cap, 2), dtype=np.uint32)
            count = C.c_int64(0)
            self.b.call("get_neighbor_list", self.h, replica, out.ctypes.data_as(_u32_p), cap,
                        C.byref(count))
            if count.value <= cap:
                return out[: count.value]
            cap = count.value

    def get_counters(self) -> Counters:
        c = Counters()
        self.b.call("get_counters", self.h, C.byref(c))
        return c

    # -- engine-only ------------------------------------------------------------------------
    def set_stream(self, cuda_stream: int):
        self.b.call("set_stream", self.h, C.c_void_p(cuda_stream))

    def sync(self):
        self.b.call("sync", self.h)

    def set_option(self, name: str, value: float):
        self.b.call("set_option", self.h, name.encode(), float(value))

    def get_option(self, name: str) -> float:
        v = C.c_double(0)
        self.b.call("get_option", self.h, name.encode(), C.byref(v))
        return v.value

    def set_positions_device(self, dptr: int):
        self.b.call("set_positions_device", self.h, C.c_void_p(dptr))

    def get_positions_device(self, dptr: int):
        self.b.call("get_positions_device", self.h, C.c_void_p(dptr))

    def set_all_positions(self, host_ptr: int):
        """host buffer [replicas][n][3] doubles (address), e.g. a pinned torch tensor's data_ptr()"""
        self.b.call("set_all_positions", self.h, C.c_void_p(host_ptr))

    def glue_update(self, bond_ff, max_distance, p_bind, p_unbind, max_glues, seed, update_index, use_box=True):
        """kinetic glue update on the device (bd_cuda.h: bd_glue_update); returns (unbound, bound, glued)"""
        a, b, c = C.c_int64(0), C.c_int64(0), C.c_int64(0)
        self.b.call("glue_update", self.h, bond_ff, 0, float(max_distance), float(p_bind), float(p_unbind), int(max_glues),
                    int(seed), int(update_index), 1 if use_box else 0, C.byref(a), C.byref(b), C.byref(c))
        return a.value, b.value, c.value

    def get_bond_pairs(self, bond_ff) -> np.ndarray:
        n = C.c_int64(0)
        self.b.call("get_bond_pairs", self.h, bond_ff, None, 0, C.byref(n))
        out = np.zeros((max(n.value, 1), 2), dtype=np.uint32)
        self.b.call("get_bond_pairs", self.h, bond_ff, out.ctypes.data_as(_u32_p), n.value, C.byref(n))
        return out[: n.value]

    def get_all_positions(self, host_ptr: int):
        self.b.call("get_all_positions", self.h, C.c_void_p(host_ptr))

    def get_all_positions_async(self, host_ptr: int):
        """enqueue the download into a pinned buffer; complete after sync()"""
        self.b.call("get_all_positions_async", self.h, C.c_void_p(host_ptr))

    def displacement_pointer(self) -> int:
        """device address of the running max displacement^2 (bits of a double); see bd_cuda.h"""
        p = C.c_void_p()
        self.b.call("displacement_pointer", self.h, C.byref(p))
        return p.value

    def set_owned(self, owned):
        """slab decomposition: 0/1 per bead, None = every bead is advanced by this engine"""
        if owned is None:
            self.b.call("set_owned", self.h, None)
            return
        m = np.ascontiguousarray(owned, dtype=np.uint8)
        assert m.shape == (self.n,)
        self.b.call("set_owned", self.h, m.ctypes.data_as(_u8_p))

    def set_owned_device(self, d_owned: int):
        """device pointer to n uint8 flags"""
        self.b.call("set_owned_device", self.h, C.c_void_p(d_owned))

    def gather_positions_indexed(self, d_index: int, count: int, d_xyz: int):
        """device pointers: uint32 original indices, xyz[count][3] doubles"""
        self.b.call("gather_positions_indexed", self.h, C.c_void_p(d_index), int(count), C.c_void_p(d_xyz))

    def scatter_positions_indexed(self, d_index: int, count: int, d_xyz: int):
        self.b.call("scatter_positions_indexed", self.h, C.c_void_p(d_index), int(count), C.c_void_p(d_xyz))

    # contact map on the device (contact_map.cc:25-91 of the interphase driver)
    def contact_update(self, distance: float, replica: int = 0):
        self.b.call("contact_update", self.h, int(replica), float(distance))

    def contact_accumulate(self, replica: int = 0) -> np.ndarray:
        """[m, 3] uint32 rows (i, j, count), i < j, sorted by (i, j)."""
        count = C.c_int64(0)
        self.b.call("contact_accumulate", self.h, int(replica), None, 0, C.byref(count))
        out = np.empty((max(count.value, 0), 3), dtype=np.uint32)
        if count.value:
            self.b.call("contact_accumulate", self.h, int(replica), out.ctypes.data_as(C.POINTER(C.c_uint32)),
                        int(count.value), C.byref(count))
        return out

    # pair search / pair distances on the resident positions (glue_simulator.cpp:35-80 of 3-sim-1kb)
    def search_pairs(self, distance: float, use_box: bool = True, replica: int = 0) -> np.ndarray:
        """[m, 2] uint32 pairs (i < j) within `distance`, sorted by (i, j)."""
        count = C.c_int64(0)
        cap = max(1024, 8 * self.n)
        while True:
            out = np.empty((cap, 2), dtype=np.uint32)
            self.b.call("search_pairs", self.h, int(replica), float(distance), int(bool(use_box)),
                        out.ctypes.data_as(_u32_p), cap, C.byref(count))
            if count.value <= cap:
                return out[: count.value]
            cap = count.value

    def pair_distances(self, pairs, min_image: bool = False, replica: int = 0) -> np.ndarray:
        p = np.ascontiguousarray(pairs, dtype=np.uint32).reshape(-1, 2)
        out = np.empty(len(p))
        self.b.call("pair_distances", self.h, int(replica), p.ctypes.data_as(_u32_p), len(p), int(bool(min_image)), _dp(out))
        return out

    # single-precision snapshots (simulation_store.cc:253-276 of 5-sim-genome)
    def snapshot_begin(self, replica: int = 0, fraction_bits: int = 16):
        self.b.call("snapshot_begin", self.h, int(replica), int(fraction_bits))

    def snapshot_wait(self) -> np.ndarray:
        """float32 [beads, 3] (all replicas concatenated when begun with replica=-1); a copy."""
        ptr, count = C.POINTER(C.c_float)(), C.c_int64(0)
        self.b.call("snapshot_wait", self.h, C.byref(ptr), C.byref(count))
        if count.value == 0:
            return np.empty((0, 3), dtype=np.float32)
        return np.ctypeslib.as_array(ptr, shape=(count.value, 3)).copy()

    def contact_clear(self, replica: int = 0):
        self.b.call("contact_clear", self.h, int(replica))

    def get_launch_count(self) -> int:
        v = C.c_int64(0)
        self.b.call("get_launch_count", self.h, C.byref(v))
        return v.value


ENGINE_LIB = os.path.join(HERE, "libbd_b200.so")
_cache = {}


def load_engine() -> Backend:
    """The product library.  Fails loudly when it has not been built: no CPU fallback exists."""
    if "bd_" not in _cache:
        lib = os.environ.get("BD_B200_LIB", ENGINE_LIB)  # developer override: A/B builds of the engine
        if lib != ENGINE_LIB:
            _cache["bd_"] = Backend(lib, "bd_")
            return _cache["bd_"]
        if not os.path.exists(ENGINE_LIB):
            raise BackendError(
                f"{ENGINE_LIB} is missing: build it with `python __graft_entry__.py build` "
                "(nvcc, sm_100a).  The engine has no CPU fallback."
            )
        _cache["bd_"] = Backend(ENGINE_LIB, "bd_")
    return _cache["bd_"]
